"""Timing probe of K5 at the BASELINE config-5 shapes (not a pytest file)."""
import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "convex-optimization_b200"))
import torch
from glprox import _lib
lib = _lib.load()
st = torch.cuda.current_stream().cuda_stream
def run(M, K, N, passes, chunk, reps=5, alias=False):
    P = torch.randn((M, K), device="cuda"); Q = torch.randn((N, K), device="cuda")
    # alias: P_lo = P_hi, same tensor work at half the DRAM bytes (the second load of a tile hits L2)
    Pl = torch.randn((M, K), device="cuda") * 1e-4 if (passes >= 2 and not alias) else P
    Ql = torch.randn((N, K), device="cuda") * 1e-4 if passes >= 2 else Q
    D = torch.empty((N, M), device="cuda", dtype=torch.float64)
    def go():
        _lib.check(lib.glp_gemm_tf32(0, st, passes, M, K, N, P.data_ptr(), Pl.data_ptr(), K, Q.data_ptr(), Ql.data_ptr(), K, D.data_ptr(), M, chunk))
    go(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): go()
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    flops = 2.0 * M * K * N * passes
    byts = M * K * 4.0 * (2 if (passes == 3 and not alias) else 1)
    print(("alias " if alias else "") + "M=%d K=%d N=%d passes=%d chunk=%d: %.3f ms  %.1f TFLOP/s (tensor work)  %.0f GB/s (P stream)" % (M, K, N, passes, chunk, ms, flops / ms / 1e9, byts / ms / 1e6), flush=True)
if len(sys.argv) > 1 and sys.argv[1] == "one":   # the launch profiled with ncu (profiles/)
    run(16384, 65536, 128, 3, 8, reps=2)
    sys.exit(0)
if len(sys.argv) > 1 and sys.argv[1] == "one4":   # the two fp16-pair launches profiled with ncu (forward and transposed shapes of C5)
    for (M, K, N) in [(16384, 65536, 128), (65536, 16384, 128)]:
        P = torch.randn((M, K), device="cuda"); Q = torch.randn((N, K), device="cuda")
        D = torch.empty((N, M), device="cuda", dtype=torch.float64)
        _lib.check(lib.glp_gemm_tf32(0, st, 4, M, K, N, P.data_ptr(), P.data_ptr(), K, Q.data_ptr(), Q.data_ptr(), K, D.data_ptr(), M, 8))
        torch.cuda.synchronize()
        del P, Q, D
    sys.exit(0)
if len(sys.argv) > 1 and sys.argv[1] == "pair":   # fp16-pair product at the solver's chunk (256 elements of K), C5 shapes; GLPROX_GEMM_PAIR selects the kernel
    for (M, K, N) in [(16384, 65536, 128), (65536, 16384, 128)]:
        run(M, K, N, 4, 4, reps=int(sys.argv[2]) if len(sys.argv) > 2 else 5)
    sys.exit(0)
if len(sys.argv) > 1 and sys.argv[1] == "alias":   # is the 3-pass product bound by the P stream from HBM?
    for (M, K, N) in [(16384, 65536, 128), (65536, 16384, 128), (16384, 65536, 256)]:
        for al in (False, True):
            run(M, K, N, 3, 8, alias=al)
        run(M, K, N, 1, 8)
    sys.exit(0)
if len(sys.argv) > 1 and sys.argv[1] == "bf16":   # 3xTF32 vs TF32 + bf16 corrections
    for (M, K, N) in [(16384, 65536, 256), (65536, 16384, 128), (16384, 8192, 256), (8192, 16384, 128)]:
        for passes in (3, 2, 1):
            run(M, K, N, passes, 8)
    sys.exit(0)
if len(sys.argv) > 1 and sys.argv[1] == "split":  # unsplit vs split-K schedule, 1-GPU and 8-GPU shard shapes
    for (M, K, N) in [(16384, 65536, 256), (65536, 16384, 128), (16384, 8192, 256), (8192, 16384, 128)]:
        for chunk in (8, -8):
            run(M, K, N, 3, chunk)
    sys.exit(0)
for (M, K, N) in [(16384, 65536, 256), (65536, 16384, 128)]:
    for passes in (1, 3):
        for chunk in (0, 8, 32):
            run(M, K, N, passes, chunk)
