"""Roofs of the hot path's kernels on this GPU (srt_measure_gather / srt_measure_issue): prints one JSON object."""
import json
import sys

sys.path.insert(0, ".")
import __graft_entry__ as g

pkg = g.load_package()
ctx = pkg.Context(0)
out = {"gather_GBps": {}, "issue_Gthreadinst_per_s": {}}
for label, n in (("l1_48KiB", 48 << 10), ("l1_64KiB", 64 << 10), ("l2_8MiB", 8 << 20), ("l2_32MiB", 32 << 20),
                 ("l2_64MiB", 64 << 20), ("hbm_1GiB", 1 << 30), ("hbm_8GiB", 8 << 30)):
    out["gather_GBps"][label] = {"independent": ctx.measure_gather(n, 0), "chase": ctx.measure_gather(n, 1)}
for mix, label in ((0, "fmul_fadd"), (1, "fmul_fadd_2fmnmx")):
    out["issue_Gthreadinst_per_s"][label] = [ctx.measure_issue(mix) for _ in range(3)]
out["nominal_issue_Gthreadinst_per_s"] = 148 * 4 * 32 * 1.965
print(json.dumps(out))
ctx.close()
