#!/usr/bin/env python
"""profiles/<name>_full_summary.csv (tools/ncu_key_metrics.py of an `ncu --set full` capture of tools/run_hot_kernels.py)
-> profiles/r02_hot_kernels_traffic.json: DRAM bytes (read + write) per launch of each hot kernel, keyed as bench.py's
kernel_rooflines, with the tree the capture was taken on.  bench.py reads this file for `roofline.traffic`.

  python tools/ncu_traffic.py profiles/r02_hot_kernels_full_summary.csv <git tree-ish> > profiles/r02_hot_kernels_traffic.json
"""
import csv
import json
import sys

# first inference repetition of tools/run_hot_kernels.py, in launch order
ORDER = [("qkv", "conv_gemm_kernel<192"), ("attention", "attention_fwd_kernel"), ("out_proj", "conv_gemm_kernel<384, 1"),
         ("ffn_fused", "ffn_fused_kernel"), ("ffn1", "conv_gemm_kernel<256"), ("ffn2", "conv_gemm_kernel<384, 2"), ("layernorm", "layernorm_kernel"),
         ("lr_expand", "lr_expand_kernel"), ("bucket_embed_add", "bucket_embed_add_kernel")]


def unit_scale(h):
    u = h[h.index("[") + 1:h.index("]")].lower() if "[" in h else "byte"
    return {"byte": 1.0, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9}.get(u, 1.0)


def main():
    rows = list(csv.reader(open(sys.argv[1])))
    hdr = rows[0]
    ir = next(i for i, h in enumerate(hdr) if h.startswith("dram_read"))
    iw = next(i for i, h in enumerate(hdr) if h.startswith("dram_write"))
    out, seen = {}, set()
    for r in rows[1:]:
        for key, pat in ORDER:
            if key not in seen and pat in r[1]:
                seen.add(key)
                out[key] = float(r[ir].replace(",", "")) * unit_scale(hdr[ir]) + float(r[iw].replace(",", "")) * unit_scale(hdr[iw])
                break
    print(json.dumps({"tree": sys.argv[2] if len(sys.argv) > 2 else None, "source": sys.argv[1],
                      "what": "dram__bytes_read.sum + dram__bytes_write.sum per launch, ncu --set full --clock-control none, "
                              "tools/run_hot_kernels.py (decoder shape of cfg2: B=64, T=1103, ragged tile list)",
                      "traffic_bytes": out}, indent=1))


if __name__ == "__main__":
    main()
