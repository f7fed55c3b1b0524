"""Host-side logic of the drop-in boundary that needs no GPU: how the reference's real argument
types -- 1-based ``pd.DataFrame``s built as search/search.py:71-72, 83-84 builds them -- are turned into upload
descriptors without copying, and ``extract_mlp`` on the model objects ``LearnedIndex.build`` leaves behind
(search/li/model.py:15-83, LearnedIndex.py:231-239)."""
import numpy as np
import pandas as pd
import pytest
import torch

from lmi_b200.index import _view_rows, _as_rows, extract_mlp


def _frames(n=500, d=64):
    rng = np.random.default_rng(1)
    arr = rng.standard_normal((n, d)).astype(np.float16)
    df = pd.DataFrame(arr)                     # search.py:83
    df.index += 1                              # search.py:84
    return arr, df


def test_view_rows_frame_is_a_view_with_1_based_ids():
    arr, df = _frames()
    rows, id_base = _view_rows(df)
    assert id_base == 1 and rows.shape == arr.shape and rows.dtype == np.float16
    assert np.shares_memory(rows, df.to_numpy())           # nothing was copied
    assert (rows == arr).all()
    es = rows.dtype.itemsize
    # a frame's block is feature-major (F-ordered view) or, on other pandas versions, row-major: both are uploadable as they lie
    assert rows.strides in ((es, es * arr.shape[0]), (es * arr.shape[1], es))
    rows_c, _ = _as_rows(df)
    assert rows_c.flags["C_CONTIGUOUS"] and (rows_c == arr).all()


def test_view_rows_drops_category_and_checks_labels():
    arr, df = _frames()
    df2 = df.copy()
    df2["category"] = 3                                    # LearnedIndex.py:140-141 drops it on a copy
    rows, id_base = _view_rows(df2)
    assert rows.shape == arr.shape and id_base == 1 and (rows == arr).all()
    assert "category" in df2.columns                       # the caller's frame is untouched
    df3 = pd.DataFrame(arr, index=np.arange(7, 7 + arr.shape[0]))
    assert _view_rows(df3)[1] == 7
    bad = pd.DataFrame(arr, index=np.arange(arr.shape[0]) * 2)
    with pytest.raises(ValueError):
        _view_rows(bad)
    rows, id_base = _view_rows(arr)                        # plain ndarray: ids are row + 1 (search.py:71-72)
    assert id_base == 1 and rows is arr or np.shares_memory(rows, arr)


class _Model(torch.nn.Module):                             # shaped like li.model.Model (model.py:15-83)
    def __init__(self, layers):
        super().__init__()
        self.layers = layers


class _NeuralNetwork:                                      # li.model.NeuralNetwork keeps the Model in .model (model.py:114-147)
    def __init__(self, m):
        self.model = m


class _LearnedIndex:                                       # li.LearnedIndex keeps the NeuralNetwork in .model (LearnedIndex.py:239)
    def __init__(self, nn):
        self.model = nn


def test_extract_mlp_on_reference_shaped_objects():
    seq = torch.nn.Sequential(torch.nn.Linear(96, 256), torch.nn.ReLU(), torch.nn.Linear(256, 128), torch.nn.ReLU(),
                              torch.nn.Linear(128, 122))          # MLP-5, model.py:45-52
    for obj in (seq, _Model(seq), _NeuralNetwork(_Model(seq)), _LearnedIndex(_NeuralNetwork(_Model(seq)))):
        w, b = extract_mlp(obj)
        assert [x.shape for x in w] == [(256, 96), (128, 256), (122, 128)]
        assert [x.shape for x in b] == [(256,), (128,), (122,)]
        assert all(x.dtype == np.float32 for x in w + b)
        assert np.array_equal(w[0], seq[0].weight.detach().numpy())
    with pytest.raises(ValueError):
        extract_mlp(torch.nn.Sequential(torch.nn.Linear(4, 4), torch.nn.Tanh(), torch.nn.Linear(4, 2)))
    with pytest.raises(ValueError):
        extract_mlp(torch.nn.Sequential(torch.nn.ReLU()))


def test_extract_mlp_on_the_real_reference_model():
    from oracle import ref_shims
    if not ref_shims.reference_available():
        pytest.skip("/root/reference not present")
    rng = np.random.default_rng(3)
    w = [rng.standard_normal((128, 96)).astype(np.float32), rng.standard_normal((17, 128)).astype(np.float32)]
    b = [rng.standard_normal(128).astype(np.float32), rng.standard_normal(17).astype(np.float32)]
    li = ref_shims.make_reference_index(w, b, "MLP")       # an unmodified li.LearnedIndex with these weights
    w2, b2 = extract_mlp(li)
    assert all(np.array_equal(x, y) for x, y in zip(w + b, w2 + b2))
