"""Dependent-chain cost inside a CUDA graph on this GPU: us per node for (a) a trivial kernel, (b) cuBLAS SGEMMs of
the PPO minibatch shapes, (c) this library's epilogue kernels. Explains where the 110 us per minibatch go."""
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
import torch  # noqa: E402

from diffusion_piano_b200.ppo import _kernels  # noqa: E402

sys.path.insert(0, str(Path(__file__).resolve().parent))
from experiments.load import load_experiments  # noqa: E402

lib = load_experiments()  # product kernels + tools/experiments (this probe's kernel is not in the product library)
dev = torch.device("cuda", 0)
S = lambda: torch.cuda.current_stream().cuda_stream  # noqa: E731


def timed(fn, n=20, reps=50):
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        for _ in range(3):
            fn()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(n):
            fn()
    g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        g.replay()
    e1.record()
    torch.cuda.synchronize()
    return 1e3 * e0.elapsed_time(e1) / (reps * n)


B, H, K = 128, 256, 329
x, w1 = torch.randn(B, K, device=dev), torch.randn(H, K, device=dev)
a, w2 = torch.randn(B, H, device=dev), torch.randn(H, H, device=dev)
y, y2 = torch.empty(B, H, device=dev), torch.empty(B, H, device=dev)
gw = torch.empty(H, H, device=dev)
one = torch.zeros(1, device=dev)
v = torch.zeros(B, device=dev)
print("trivial kernel (add_bias, 128 elements)      %.2f us/node" % timed(lambda: lib.ppo_add_bias(v.data_ptr(), one.data_ptr(), v.data_ptr(), B, S())))
big = torch.zeros(128 * 256, device=dev)
for mode, name in ((0, "plain"), (1, "PDL"), (2, "PDL + early trigger")):
    print("probe 128 elements, %-20s        %.2f us/node" % (name, timed(lambda: lib.ppo_pdl_probe(v.data_ptr(), B, mode, S()))))
    print("probe 32768 elements, %-20s      %.2f us/node" % (name, timed(lambda: lib.ppo_pdl_probe(big.data_ptr(), 128 * 256, mode, S()))))
print("mm fwd   [128,329] x [256,329]^T             %.2f us/node" % timed(lambda: torch.mm(x, w1.t(), out=y)))
print("mm fwd   [128,256] x [256,256]^T (dependent) %.2f us/node" % timed(lambda: (torch.mm(a, w2.t(), out=y), torch.mm(y, w2.t(), out=a))) + "  (per pair / 2 below)")
print("mm dA    [128,256] x [256,256]               %.2f us/node" % timed(lambda: torch.mm(a, w2, out=y2)))
print("mm dW    [256,128] x [128,256]               %.2f us/node" % timed(lambda: torch.mm(a.t(), y, out=gw)))
print("ppo_gemm_nt [128,329] x [256,329]^T          %.2f us/node" % timed(lambda: lib.ppo_gemm_nt(x.data_ptr(), w1.data_ptr(), y.data_ptr(), B, H, K, S())))
print("ppo_gemm_nt [128,256] x [256,256]^T          %.2f us/node" % timed(lambda: lib.ppo_gemm_nt(a.data_ptr(), w2.data_ptr(), y.data_ptr(), B, H, H, S())))
print("ppo_gemm_nn [128,256] x [256,256]            %.2f us/node" % timed(lambda: lib.ppo_gemm_nn(a.data_ptr(), w2.data_ptr(), y2.data_ptr(), B, H, H, S())))
print("ppo_gemm_tn [256,128] x [128,256]            %.2f us/node" % timed(lambda: lib.ppo_gemm_tn(a.data_ptr(), y.data_ptr(), gw.data_ptr(), B, H, H, S())))
gw1 = torch.empty(H, K, device=dev)
print("ppo_gemm_tn [256,128] x [128,329]            %.2f us/node" % timed(lambda: lib.ppo_gemm_tn(a.data_ptr(), x.data_ptr(), gw1.data_ptr(), B, H, K, S())))
bias, gam = torch.zeros(H, device=dev), torch.ones(H, device=dev)
out, st = torch.empty(B, H, device=dev), torch.empty(B, 2, device=dev)
print("relu_ln_fwd 128x256                          %.2f us/node" % timed(lambda: lib.ppo_relu_ln_fwd(y.data_ptr(), bias.data_ptr(), gam.data_ptr(), bias.data_ptr(), out.data_ptr(), None, st.data_ptr(), B, H, 0.0, None, None, None, 0, S())))
dy, db, dg, dbe = torch.empty(B, H, device=dev), torch.empty(H, device=dev), torch.empty(H, device=dev), torch.empty(H, device=dev)
print("relu_ln_bwd_dy 128x256                       %.2f us/node" % timed(lambda: lib.ppo_relu_ln_bwd_dy(out.data_ptr(), y.data_ptr(), st.data_ptr(), gam.data_ptr(), None, 0.0, dy.data_ptr(), B, H, S())))
print("relu_ln_colsums 128x256                      %.2f us/node" % timed(lambda: lib.ppo_relu_ln_colsums(out.data_ptr(), y.data_ptr(), st.data_ptr(), None, 0.0, dy.data_ptr(), db.data_ptr(), dg.data_ptr(), dbe.data_ptr(), B, H, S())))
for tf32 in (True,):
    torch.backends.cuda.matmul.allow_tf32 = tf32
    print("mm fwd tf32 [128,329] x [256,329]^T          %.2f us/node" % timed(lambda: torch.mm(x, w1.t(), out=y)))
    print("mm dA  tf32 [128,256] x [256,256]            %.2f us/node" % timed(lambda: torch.mm(a, w2, out=y2)))
