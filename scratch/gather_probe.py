"""Random 64 B gather rates of this GPU (srt_measure_gather): the denominators DESIGN.md quotes for the soup."""
import json, sys
sys.path.insert(0, '.')
import __graft_entry__ as g
pkg = g.load_package()
ctx = pkg.Context(0)
out = {}
for gb in (0.064, 1.1, 8.0):
    for chase in (0, 1):
        v = [ctx.measure_gather(int(gb * 1e9), chase) for _ in range(3)]
        out["%.3f GB array, %s" % (gb, "dependent chain" if chase else "independent")] = round(max(v), 1)
print(json.dumps(out, indent=1))
