"""TF32 and fp32-FFMA dense peaks of this GPU, measured the way MEASURED_PEAKS.json was made (torch.matmul on
8192^3, 2*N^3 flops, best of 10 with CUDA events, plus a 4 s back-to-back sustained figure), and the HBM copy rate
for reference.  Writes profiles/r2_peaks_tf32_ffma.json, which bench.py uses as the roofline denominator of the
tensor-bound kernels in the fp32 / tf32 / tf32x3 math modes (the driver's file only has bf16).

  python scripts/measure_peaks.py
"""
import json
import os
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
N = 8192


def rate(fn, flops, seconds=4.0):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    best = 1e9
    for _ in range(10):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0, n = time.perf_counter(), 0
    e0.record()
    while time.perf_counter() - t0 < seconds:
        for _ in range(10):
            fn()
        n += 10
        torch.cuda.synchronize()
    e1.record()
    torch.cuda.synchronize()
    return flops / (best * 1e-3) / 1e12, flops * n / (e0.elapsed_time(e1) * 1e-3) / 1e12


def main():
    a = torch.randn(N, N, device="cuda")
    b = torch.randn(N, N, device="cuda")
    c = torch.empty(N, N, device="cuda")
    flops = 2.0 * N ** 3
    torch.backends.cuda.matmul.allow_tf32 = True
    tf32, tf32_s = rate(lambda: torch.matmul(a, b, out=c), flops)
    torch.backends.cuda.matmul.allow_tf32 = False
    ffma, ffma_s = rate(lambda: torch.matmul(a, b, out=c), flops, seconds=2.0)
    ab, bb = a.bfloat16(), b.bfloat16()
    cb = torch.empty(N, N, device="cuda", dtype=torch.bfloat16)
    bf16, bf16_s = rate(lambda: torch.matmul(ab, bb, out=cb), flops)
    src = torch.empty(1 << 30, dtype=torch.bfloat16, device="cuda")
    dst = torch.empty_like(src)
    best = 1e9
    for _ in range(10):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        dst.copy_(src)
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    out = {"tf32_tflops": tf32, "tf32_tflops_sustained": tf32_s, "fp32_ffma_tflops": ffma, "fp32_ffma_tflops_sustained": ffma_s,
           "bf16_tflops_here": bf16, "bf16_tflops_sustained_here": bf16_s, "hbm_gbs_here": 2 * src.numel() * 2 / (best * 1e-3) / 1e9,
           "gpu_name": torch.cuda.get_device_name(0), "torch": torch.__version__,
           "how": "torch.matmul fp32 8192^3 with allow_tf32 = True (cuBLAS TF32 tensor cores) and False (cuBLAS fp32: FFMA), "
                  "2*N^3 flops, best of 10 (burst) and back to back for 4 s / 2 s (sustained), CUDA events; bf16 and the "
                  "1 Gi-element bf16 copy repeated here as a cross-check of MEASURED_PEAKS.json on this box"}
    path = os.path.join(ROOT, "gpurun_out", "r2_peaks_tf32_ffma.json")
    os.makedirs(os.path.dirname(path), exist_ok=True)
    with open(path, "w") as handle:
        json.dump(out, handle, indent=1)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
