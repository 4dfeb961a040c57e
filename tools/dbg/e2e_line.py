import json, sys
d = json.loads(sys.stdin.read().strip().splitlines()[-1])
e = d["e2e"]
print("value %.2f e2e %.3f Gbit/s = %.2f M frames/s  bound %.2f M (W %.1f GB/s, R %.2f Gv/s) frac %.2f" % (
    d["value"], e["value"], e["frames_per_s"] / 1e6, e["bound"]["bound_frames_per_s"] / 1e6, e["bound"]["h2d_GBps_sum"],
    e["bound"]["host_narrow_Gvalues_per_s_sum"], e["bound"]["frac_of_bound"]))
