"""Where the end-to-end call spends its time outside the device loop: the one-shot update_div is
session create (allocation) + upload (H2D) + run + download (D2H) + destroy; this times each of them
with pinned host buffers.   python scripts/e2e_breakdown.py M N K [steps]"""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import nmf_gpu_b200.api as api

M, N, K = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
steps = int(sys.argv[4]) if len(sys.argv) > 4 else 20
api.lib()
torch.cuda.init(); torch.cuda.synchronize()
t = time.perf_counter()
Xh = torch.empty((N, M), dtype=torch.float32, pin_memory=True).uniform_(0, 1)
Wh = torch.empty((K, M), dtype=torch.float32, pin_memory=True).uniform_(0, 1)
Hh = torch.empty((N, K), dtype=torch.float32, pin_memory=True).uniform_(0, 1)
host_s = time.perf_counter() - t
Xn, Wn, Hn = Xh.numpy().T, Wh.numpy().T, Hh.numpy().T
out = {"M": M, "N": N, "K": K, "steps": steps, "pinned_alloc_and_fill_s": host_s}
for rep in range(2):
    marks = [time.perf_counter()]
    s = api.Session(M, N, K); torch.cuda.synchronize(); marks.append(time.perf_counter())
    s.upload(Wn, Hn, Xn); torch.cuda.synchronize(); marks.append(time.perf_counter())
    r = s.run(0.0, steps); torch.cuda.synchronize(); marks.append(time.perf_counter())
    s.download(); torch.cuda.synchronize(); marks.append(time.perf_counter())
    s.close(); torch.cuda.synchronize(); marks.append(time.perf_counter())
    d = [b - a for a, b in zip(marks, marks[1:])]
    out[f"rep{rep}"] = {"create_s": d[0], "upload_s": d[1], "upload_GBps": 4.0 * M * N / d[1] / 1e9, "run_s": d[2], "loop_device_ms": r.device_ms,
                       "download_s": d[3], "destroy_s": d[4], "total_s": marks[-1] - marks[0]}
t = time.perf_counter()
r = api.update_div(Wn, Hn, Xn, 0.0, steps); torch.cuda.synchronize()
out["one_shot_update_div_s"] = time.perf_counter() - t
out["one_shot_loop_device_ms"] = r.device_ms
print(json.dumps(out))
