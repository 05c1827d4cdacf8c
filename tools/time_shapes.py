"""GPU-box diagnostic: CUDA-event time of every GEMM call shape of one cfg-2 conditioner (B = 65536),
through the C ABI with the row pitches the library uses (16 / 128 / 324 floats)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from vcnf_b200 import _lib
from vcnf_b200._lib import ptr, stream

B = 65536
dev = "cuda"
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def pad(rows, cols, pitch):
    return torch.randn(rows, pitch, device=dev)[:, :cols]


def timeit(fn, reps=10):
    ts = []
    for _ in range(reps):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    ts.sort()
    return ts[len(ts) // 2]


def lin(A, W, sw_n, sw_k, bias, res, mask, C, N, K, relu):
    _lib.call("vcnf_linear_tf32x3", ptr(A), A.stride(0), ptr(W), sw_n, sw_k, ptr(bias), ptr(res), res.stride(0) if res is not None else 0,
              ptr(mask), mask.stride(0) if mask is not None else 0, ptr(C), C.stride(0), B, N, K, int(relu), stream())


x16, h, h2, p324 = pad(B, 15, 16), pad(B, 128, 128), pad(B, 128, 128), pad(B, 322, 324)
W_in, W_h, W_out = torch.randn(128, 15, device=dev), torch.randn(128, 128, device=dev), torch.randn(322, 128, device=dev)
b128, b322 = torch.randn(128, device=dev), torch.randn(322, device=dev)
out128, g16 = pad(B, 128, 128), pad(B, 15, 16)
rows = []
rows.append(("fwd  input  15->128 (+bias)", 16, timeit(lambda: lin(x16, W_in, 15, 1, b128, None, None, out128, 128, 15, False)), 4 * (16 + 128)))
rows.append(("fwd  hidden 128->128 relu+bias", 32, timeit(lambda: lin(h, W_h, 128, 1, b128, None, None, out128, 128, 128, True)), 4 * 256))
rows.append(("fwd  hidden 128->128 relu+bias+res", 32, timeit(lambda: lin(h, W_h, 128, 1, b128, h2, None, out128, 128, 128, True)), 4 * 384))
rows.append(("fwd  head   128->322 (+bias)", 16, timeit(lambda: lin(h, W_out, 128, 1, b322, None, None, p324, 322, 128, False)), 4 * (128 + 322)))
rows.append(("dgrad hidden (W^T, relu mask)", 64, timeit(lambda: lin(h, W_h, 1, 128, None, None, h2, out128, 128, 128, False)), 4 * 384))
rows.append(("dgrad input  128->15", 16, timeit(lambda: lin(h, W_in, 1, 15, None, None, None, g16, 15, 128, False)), 4 * (128 + 16)))


def head_dgrad():
    first = True
    for n0 in range(0, 322, 128):
        n1 = min(322, n0 + 128)
        _lib.call("vcnf_linear_tf32x3", ptr(p324) + 4 * n0, 324, ptr(W_out) + 4 * n0 * 128, 1, 128, None,
                  None if first else ptr(out128), 128, None, 0, ptr(out128), 128, B, 128, n1 - n0, 0, stream())
        first = False


rows.append(("dgrad head 322->128 (3 K-slices)", 16, timeit(head_dgrad), 4 * (322 + 128 * 5)))
gW, gb = torch.empty(128, 128, device=dev), torch.empty(322, device=dev)
gW_in, gW_out = torch.empty(128, 15, device=dev), torch.empty(322, 128, device=dev)
ws = torch.empty(_lib.query("vcnf_linear_wgrad_workspace", B, 322, 128), dtype=torch.uint8, device=dev)


def wg(G, X, out, N, K, relu):
    _lib.call("vcnf_linear_wgrad_tf32x3", ptr(G), G.stride(0), ptr(X), X.stride(0), ptr(out), ptr(gb), ptr(ws), B, N, K, int(relu), stream())


rows.append(("wgrad input  (N=128, K=15)", 16, timeit(lambda: wg(h, x16, gW_in, 128, 15, False)), 4 * (128 + 16)))
rows.append(("wgrad hidden (N=128, K=128, relu)", 64, timeit(lambda: wg(h, h2, gW, 128, 128, True)), 4 * 256))
rows.append(("wgrad head   (N=322, K=128)", 16, timeit(lambda: wg(p324, h, gW_out, 322, 128, False)), 4 * (322 + 128)))
tot = 0.0
for name, n, us, bps in rows:
    tot += n * us
    print("%-38s %7.1f us x %3d = %6.2f ms   %5.0f GB/s" % (name, us, n, n * us / 1e3, bps * B / us / 1e3))
print("sum over a step: %.2f ms" % (tot / 1e3))
