"""TF32 tensor-core rates to put TF32 fractions against ("TF32 peak to be measured by the builder with the
same method before any TF32 fraction is quoted"; MEASURED_PEAKS.json holds only BF16): torch.matmul FP32 with
allow_tf32, 8192^3 and 4096^3, best of 10 (burst) and back to back for ~3 s (sustained), CUDA events.
Writes profiles/tf32_peaks.json (read by bench.py --precision tf32 when present).  Run on the GPU box."""
import json, os, sys, time
import torch

torch.backends.cuda.matmul.allow_tf32 = True
out = {}
for N in (8192, 4096):
    a = torch.randn(N, N, device="cuda", dtype=torch.float32)
    b = torch.randn(N, N, device="cuda", dtype=torch.float32)
    for _ in range(5):
        a @ b
    torch.cuda.synchronize()
    best = 1e9
    for _ in range(10):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); a @ b; e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    t0 = time.perf_counter()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    n = 0
    while time.perf_counter() - t0 < 3.0:
        for _ in range(20):
            a @ b
        n += 20
        torch.cuda.synchronize()
    e1.record(); torch.cuda.synchronize()
    sus = e0.elapsed_time(e1) / n
    out["n%d" % N] = {"burst_tflops": 2.0 * N ** 3 / best / 1e9, "sustained_tflops": 2.0 * N ** 3 / sus / 1e9}
    print(N, out["n%d" % N], flush=True)
out["tf32_tflops"] = out["n8192"]["burst_tflops"]
out["tf32_tflops_sustained"] = out["n8192"]["sustained_tflops"]
out["how"] = "torch.matmul fp32 with allow_tf32, 8192^3: best of 10 (burst), back to back for 3 s (sustained), CUDA events"
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
with open(os.path.join(root, "gpurun_out", "tf32_peaks.json"), "w") as f:
    json.dump(out, f, indent=1)
print(json.dumps(out))
