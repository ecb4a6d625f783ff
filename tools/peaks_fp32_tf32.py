""""measure fp32-FMA and TF32 matmul peaks on the box the same way MEASURED_PEAKS.json measured bf16
(best-of-10 torch.matmul 8192^3)". Writes gpurun_out/r02_peaks_fp32_tf32.json (copied to profiles/)."""
import json, sys
import torch
dev = torch.device("cuda:0")
N = 8192
res = {"how": "torch.matmul 8192^3, 2*N^3 flops, best of 10, CUDA events; library GEMMs used as yardsticks only", "gpu": torch.cuda.get_device_name(0),
       "torch": torch.__version__}
def best(fn, n=10):
    fn(); torch.cuda.synchronize()
    b = 1e9
    for _ in range(n):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        b = min(b, e0.elapsed_time(e1))
    return b
a = torch.randn(N, N, device=dev); b = torch.randn(N, N, device=dev)
torch.backends.cuda.matmul.allow_tf32 = False
res["fp32_ffma_tflops"] = 2.0 * N ** 3 / (best(lambda: a @ b) / 1e3) / 1e12
torch.backends.cuda.matmul.allow_tf32 = True
res["tf32_tflops"] = 2.0 * N ** 3 / (best(lambda: a @ b) / 1e3) / 1e12
torch.backends.cuda.matmul.allow_tf32 = False
ah, bh = a.bfloat16(), b.bfloat16()
res["bf16_tflops"] = 2.0 * N ** 3 / (best(lambda: ah @ bh) / 1e3) / 1e12
x = torch.empty(1 << 30, dtype=torch.bfloat16, device=dev); y = torch.empty_like(x)
res["hbm_copy_gbs"] = 2.0 * x.numel() * 2 / (best(lambda: y.copy_(x)) / 1e3) / 1e9
res["split_bf16_equivalent_tflops"] = res["bf16_tflops"] / 3.0
res["note"] = ("the path's matrix products run as 3 bf16 MMAs per fp32 product (split-bf16, fp32 accumulate): its ceiling in "
               "algorithmic fp32 FLOPs is bf16 / 3; compare with what cuBLAS reaches in plain fp32 (FFMA) and TF32")
json.dump(res, open("gpurun_out/r02_peaks_fp32_tf32.json", "w"), indent=1)
print(json.dumps(res))
