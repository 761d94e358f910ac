"""Dev: print the numbers of a bench.py JSON line (file may hold other lines, e.g. NCCL's banner)."""
import json, sys
for path in sys.argv[1:]:
    for l in open(path):
        if not l.startswith("{"):
            continue
        d = json.loads(l)
        if d.get("impl") == "reference":
            print(path, "reference arm:", d.get("value"), d.get("unit"), {k: v.get("value") for k, v in d.get("configs", {}).items()})
            continue
        r = d["roofline"]
        print("%s N=%d headline %.4e ms %.2f e2e %.4e frac %.3f pipe_frac %s gather %s clocks %s" % (
            path, d["n_gpus"], d["value"], d["ms_per_step"], d["e2e"]["value"], r["frac"], r.get("pipe_frac"), d.get("gather_check"), d.get("clocks")))
        f = d.get("full_trajectory_mode")
        if f: print("   FULL: %.2f ms %.0f GB/s hbm_frac %.3f" % (f.get("ms_per_step", f.get("ms", 0)), f["hbm_write_gbs"], f["hbm_frac"]))
        t = d.get("e2e_trazar_full")
        if t: print("   e2e_trazar_full %.3e (%.2f s, %.1f GB/s)" % (t["value"], t["seconds"], t["d2h_gbs"]))
        for k, v in d.get("configs", {}).items():
            rr = v["roofline"]
            print("   %-22s %.4e ms %.1f e2e %.4e frac %s pipe_frac %s kernel_ms %s gather %s" % (
                k, v["value"], v["ms_per_step"], v["e2e"]["value"], rr.get("frac"), rr.get("pipe_frac"), rr.get("kernel_ms"), v.get("gather_check")))
        c = d.get("cpu_baseline")
        if c: print("   cpu_baseline %.3e %s cores %d" % (c["value"], c["kind"], c["cores"]))
