import json, sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print("value %.4g DoF/s  ms/step %.2f  e2e %.4g (%.1f ms) its %s l2 %.4e launches %s" % (d["value"], d["ms_per_step"], d["e2e"]["value"], d["e2e"]["ms_per_step"], d["cg_iterations"], d["l2_error"], d["gpu_launches"]))
for k,v in d["matvec"].items():
    print("  matvec %s: %.4g DoF/s  %.3f ms  hbm frac %.3f  flop frac %.3f" % (k, v["dofs_per_s"], v["ms"], v["frac_hbm"], v["frac_cuda_core_peak"]))
for k,v in d["kernels_finest_level"].items():
    if v: print("  %-22s n=%d %.3f ms/launch share %.3f  %.0f GB/s frac %.3f" % (k, v["launches"], v["ms_per_launch"], v["share_of_step"], v["achieved_gbs"], v["frac_hbm"]))
print("  clocks", d["clocks"], "cpu", d.get("cpu_baseline"))
