"""Print value / frame ms / selected kernel times of gpurun_out/var_*.json bench lines."""
import glob, json, sys
keys = sys.argv[1:] or ["k_bounce_begin", "k_seg_exits", "k_entry_camera", "k_query_encode", "k_query_mlp"]
for f in sorted(glob.glob("gpurun_out/var_*.json")):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
    except Exception as e:  # noqa: BLE001
        print(f, "ERR", e)
        continue
    k = d.get("kernels", {})
    print(f"{f[15:-5]:>22}  {d['value']:8.1f} Mrays/s  {d['ms_per_step']:.4f} ms  e2e {d['e2e']['value']:8.1f}  " +
          "  ".join(f"{n[2:]} {k[n]['ms_per_step']:.4f}" for n in keys if n in k))
