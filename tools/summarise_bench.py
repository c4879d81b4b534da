#!/usr/bin/env python3
"""Print the headline fields of bench.py JSON lines (diagnostic)."""
import json
import sys
for f in sys.argv[1:]:
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
    except Exception as e:
        print(f, "unreadable:", e)
        continue
    print("==", f, "N =", d["n_gpus"], "value", round(d["value"]), "roofline", round(d["roofline"]["frac"], 3),
          "single", round(d["roofline"]["single_launch"]["frac"], 3), "e2e_kernel_us", round(d["roofline"]["e2e_kernel"]["us"], 2))
    for k in ("e2e", "e2e_ipc", "e2e_shm_wire", "e2e_shm_wire_in_place", "e2e_replay", "e2e_replay_cache_host"):
        v = d.get(k)
        if v:
            print("  %-22s %9.0f img/s  %.3f of PCIe  check=%s recv=%s send=%s" % (k, v["value"], v["frac_of_pcie_h2d"], v["e2e_check"],
                                                                                 v.get("receivers"), v.get("senders")))
    for k, v in sorted(d.get("configs", {}).items()):
        print("  %-22s %9.0f img/s  %.3f of PCIe  check=%s %s" % (k, v["value"], v["frac_of_pcie_h2d"], v["e2e_check"],
                                                                  {a: round(b, 1) for a, b in v.get("kernels", {}).items()}))
    h = d["host"]
    print("  host: cpus", h["cpus"], "pcie", [round(x, 1) for x in h["pcie_h2d_gbs_per_gpu_concurrent"]],
          "memcpy", [round(x, 1) for x in h["host_memcpy_gbs_per_rank_concurrent"]], "socket_path", round(h["socket_path_images_per_s"]),
          h["thread_split"])
    print("  cpu_baseline:", d.get("cpu_baseline"))
