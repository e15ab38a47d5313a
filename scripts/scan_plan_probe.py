"""Developer probe: config-3 scan time (10 M x 512 fp16, Q = 10 000, k = 100) under plan knobs, one subprocess per setting.
   python scripts/scan_plan_probe.py "EVR_SCAN_RING=20" "EVR_SCAN_RING=20 EVR_SCAN_ROTATE=1" ..."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def one(q=10_000, n=10_000_000, iters=8):
    import torch
    import __graft_entry__ as ge
    import bench
    pkg = ge.load_package()
    dev = torch.device("cuda:0")
    store = pkg.search.FrameStore(bench.make_store(torch, dev, n, seed=2))
    g = torch.Generator().manual_seed(3)
    qq = torch.randn((q, 512), generator=g)
    qq = (qq / qq.norm(dim=-1, keepdim=True)).to(dev).half()
    for _ in range(3):
        sc, idx = store.search(qq, 100)
    times = []
    for _ in range(iters):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        sc, idx = store.search(qq, 100)
        e1.record()
        torch.cuda.synchronize()
        times.append(e0.elapsed_time(e1))
    times.sort()
    print(json.dumps({"env": os.environ.get("EVR_PROBE_LABEL", ""), "Q": q, "N": n, "median_ms": round(times[len(times) // 2], 3),
                      "min_ms": round(times[0], 3), "qps": round(q / times[len(times) // 2] * 1e3),
                      "checksum": float(sc.double().sum()), "idx_sum": int(idx.sum())}), flush=True)


if __name__ == "__main__":
    if os.environ.get("EVR_PROBE_CHILD"):
        one(int(os.environ.get("EVR_PROBE_Q", "10000")), int(os.environ.get("EVR_PROBE_N", "10000000")),
            int(os.environ.get("EVR_PROBE_ITERS", "8")))
    else:
        for setting in sys.argv[1:] or [""]:
            env = dict(os.environ, EVR_PROBE_CHILD="1", EVR_PROBE_LABEL=setting)
            for kv in setting.split():
                k, v = kv.split("=")
                env[k] = v
            subprocess.run([sys.executable, __file__], env=env)
