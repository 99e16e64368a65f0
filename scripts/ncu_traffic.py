"""DRAM bytes per launch of every kernel of one training step, from an ncu metrics pass
    ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --csv --log-file X.csv \
        env B200IQ_GRAPH=0 python bench.py --steps 2 --warmup 3 --no-roofline --no-cpu-baseline --no-secondary [--workload lstm]
  python scripts/ncu_traffic.py transformer gpurun_out/X.csv [lstm gpurun_out/Y.csv] > profiles/r02_kernel_traffic.json
bench.py reads the JSON for `roofline.traffic` (keys = the names of ops.KERNEL_TRACE / GEMM entries)."""
import collections
import csv
import json
import re
import sys

# kernel function name (regex) -> the C entry names bench.py's trace uses
MAP = [
    (r"gemm_f16_kernel", ["gemm_f16_kernel"]),
    (r"attention_fwd", ["attention_fwd"]),
    (r"attention_bwd", ["attention_bwd_f16", "attention_bwd"]),
    (r"ln_fwd_warp_kernel|layernorm_fwd", ["layernorm_fwd_f16", "layernorm_fwd", "add_layernorm_fwd"]),
    (r"ln_bwd_warp_kernel|layernorm_bwd", ["layernorm_bwd_f16", "layernorm_bwd"]),
    (r"colsum_f16", ["colsum_add_f16"]),
    (r"colsum4_kernel<\(bool\)1>|colsum4_kernel<true>|relu_bwd_colsum", ["relu_bwd_colsum_f16", "relu_bwd_colsum"]),
    (r"colsum4_kernel<\(bool\)0>|colsum4_kernel<false>|colsum_kernel", ["colsum_add"]),
    (r"ce_rows_f16", ["ce_rows_f16"]),
    (r"ce_aux", ["ce_aux"]),
    (r"adam_kernel", ["adam_step"]),
    (r"sumsq_kernel", ["grad_sumsq"]),
    (r"split_f16", ["split_f16"]),
    (r"lstm_seq_fwd", ["lstm_seq_fwd"]),
    (r"lstm_seq_bwd", ["lstm_seq_bwd"]),
]


def one_step(path):
    rows = [r for r in csv.reader(open(path)) if len(r) > 5]
    hdr = [i for i, r in enumerate(rows) if r[0] == "ID"][0]
    h, data = rows[hdr], rows[hdr + 1:]
    ii, ki, mi, vi, ui = h.index("ID"), h.index("Kernel Name"), h.index("Metric Name"), h.index("Metric Value"), h.index("Metric Unit")
    launches = collections.OrderedDict()
    for r in data:
        d = launches.setdefault(r[ii], {"k": r[ki]})
        v = float(r[vi].replace(",", ""))
        u = r[ui]
        if r[mi].startswith("gpu__time"):
            d["us"] = v / 1e3 if u in ("ns", "nsecond") else (v * 1e3 if u in ("ms", "msecond") else v)
        else:
            mult = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(u, 1)
            d["bytes"] = d.get("bytes", 0.0) + v * mult
    ls = list(launches.values())
    adam = [i for i, d in enumerate(ls) if "adam_kernel" in d["k"]]
    return ls[adam[-2] + 1:adam[-1] + 1]


def main():
    out = {"source": "ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum over one eager step: " + ", ".join(sys.argv[2::2])}
    for wl, path in zip(sys.argv[1::2], sys.argv[2::2]):
        step = one_step(path)
        res = {}
        for rx, names in MAP:
            sel = [d for d in step if re.search(rx, d["k"])]
            if not sel:
                continue
            e = {"launches": len(sel), "dram_bytes_per_launch": sum(d.get("bytes", 0.0) for d in sel) / len(sel),
                 "us_per_launch": sum(d.get("us", 0.0) for d in sel) / len(sel)}
            for n in names:
                res[n] = e
        res["_step"] = {"launches": len(step), "us_serialised": sum(d.get("us", 0.0) for d in step), "dram_bytes": sum(d.get("bytes", 0.0) for d in step)}
        out[wl] = res
    json.dump(out, sys.stdout, indent=1)


if __name__ == "__main__":
    main()
