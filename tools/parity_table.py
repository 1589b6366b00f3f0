"""gpurun_out/parity_errors_<library>.json (written by tests/test_parity_fullsize_gpu.py) -> a markdown table.
  python tools/parity_table.py gpurun_out/parity_errors_libb200mt.json gpurun_out/parity_errors_libb200mt_ieee.json > profiles/parity_errors_r02.md"""
import json
import sys


def main():
    docs = [json.load(open(p)) for p in sys.argv[1:]]
    print("# Element-wise error against the reference's own kernels at the BASELINE sizes (round 2)\n")
    print("Written by `tests/test_parity_fullsize_gpu.py` on a B200; reference = `oracle/_ref/amp_C_ref.so` (the reference's unmodified "
          "`csrc/multi_tensor_*.cu`, `--use_fast_math`).  `max` / `p99.9` are the maximum and the 99.9th percentile of |ours - ref| / |ref| "
          "over all elements (fp64).  A large `max` with a tiny `p99.9` is a handful of elements that land next to zero (p_new = p - step, "
          "m = b1*m + (1-b1)*g): the row `p_over_p_plus_step` measures the same error against |p| + |applied step|, which is the bound the "
          "tests assert (1e-6 Adam, 1e-5 LAMB).  Libraries: `libb200mt` = the shipped build (approximate div/sqrt like the reference), "
          "`libb200mt_ieee` = `-DB200MT_APPROX_MATH=0`.\n")
    libs = [d["library"] for d in docs]
    print("| case | quantity | " + " | ".join("%s max | %s p99.9" % (l, l) for l in libs) + " |")
    print("|---|---|" + "---:|---:|" * len(libs))
    cases = list(docs[0]["rows"])
    for c in cases:
        for q in docs[0]["rows"][c]:
            cells = []
            for d in docs:
                r = d["rows"].get(c, {}).get(q)
                if r is None:
                    cells += ["-", "-"]
                elif "count" in r:
                    cells += ["%d of %d" % (r["count"], r["of"]), ""]
                else:
                    cells += ["%.2e" % r["max_rel"], "%.2e" % r["p99.9_rel"]]
            print("| %s | %s | %s |" % (c, q, " | ".join(cells)))


if __name__ == "__main__":
    main()
