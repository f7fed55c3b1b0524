"""lmi_b200 -- B200-native (sm_100a) query-time hot path of the LMI learned index.

Navigation (fp32 MLP) -> top-bucket selection -> exact clip768 rescoring -> top-k, behind the
reference's ``LearnedIndex.search`` / ``search_single`` / ``predict_proba`` calling convention.
The compute lives in ``csrc/`` (hand-written CUDA, C-ABI in ``include/lmi_b200.h``); this
package is the thin host-side mirror of the reference interface.  There is no CPU fallback:
importing ``lmi_b200.index`` fails loudly when ``liblmi_b200.so`` is missing.
"""
__version__ = "0.1.0"
