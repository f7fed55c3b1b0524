"""Summarise an .ncu-rep (raw page + source page) into text: python scripts/ncu_summary.py rep [--src]"""
import csv, subprocess, sys, io
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
want = ['Kernel Name', 'gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'gpu__dram_throughput.avg.pct',
        'lts__throughput.avg.pct', 'lts__t_sector_hit_rate.pct', 'lts__t_bytes.sum ', 'sm__pipe_tensor_cycles_active.avg', 'sm__warps_active.avg',
        'launch__registers_per_thread', 'sm__throughput.avg.pct', 'launch__grid_size', 'launch__block_size', 'smsp__cycles_active.avg',
        'sm__inst_executed_pipe_tensor', 'l1tex__data_bank_conflicts', 'smsp__inst_executed.sum ', 'sm__cycles_elapsed.max']
for vals in rows[2:]:
    for h, u, v in zip(hdr, units, vals):
        if any(w in h + ' ' for w in want):
            print(f"{h} [{u}] = {v}")
if '--src' in sys.argv:
    src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(src)))
    hdr = rows[1]; data = [r for r in rows[2:] if len(r) > 10]
    i_src, i_s, i_ex = hdr.index('Source'), hdr.index('# Samples'), hdr.index('Instructions Executed')
    tot = sum(int(r[i_s] or 0) for r in data)
    print(f"# source page: {len(data)} SASS instructions, {tot} stall samples; regions of 60 instructions: first_index samples max_executed top_opcodes")
    for b in range(0, len(data), 60):
        seg = data[b:b + 60]
        s = sum(int(r[i_s] or 0) for r in seg); e = max(int(r[i_ex] or 0) for r in seg)
        ops = {}
        for r in seg:
            t = r[i_src].split()
            op = t[1] if t[0].startswith('@') else t[0]
            ops[op] = ops.get(op, 0) + 1
        print(b, s, e, sorted(ops.items(), key=lambda t: -t[1])[:5])
    print("# top 20 instructions by samples")
    for r in sorted(data, key=lambda r: -int(r[i_s] or 0))[:20]:
        print(r[i_s], r[i_ex], r[i_src][:100])
