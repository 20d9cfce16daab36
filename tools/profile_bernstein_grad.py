"""Time the Bernstein reverse-mode kernels (csrc/bernstein_grad.cu) beside the forward kernel.
    python tools/profile_bernstein_grad.py [order] [log2 n]"""
import sys

import torch

sys.path.insert(0, ".")
from normalizing_flows_b200 import ops  # noqa: E402

order = int(sys.argv[1]) if len(sys.argv) > 1 else 9
n = 1 << (int(sys.argv[2]) if len(sys.argv) > 2 else 24)
raw = torch.randn(order, device="cuda")
th = ops.bernstein_constrain_thetas(raw)
x = torch.rand(n, device="cuda")
yb, lb = torch.randn(n, device="cuda"), torch.randn(n, device="cuda")


def timeit(fn, reps=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps


t_f = timeit(lambda: ops.bernstein_forward_and_log_det(th, x))
t_b = timeit(lambda: ops.bernstein_forward_and_log_det_vjp(th, x, yb, lb))
print(f"order {order} n {n}: forward {t_f*1e3:.1f} us ({12*n/t_f/1e6:.0f} GB/s), "
      f"vjp shared {t_b*1e3:.1f} us ({16*n/t_b/1e6:.0f} GB/s algorithmic)")
nr = n >> 3
thr = ops.bernstein_constrain_thetas(torch.randn(nr, order, device="cuda"))
xr, ybr, lbr = x[:nr].contiguous(), yb[:nr].contiguous(), lb[:nr].contiguous()
t_fr = timeit(lambda: ops.bernstein_forward_and_log_det(thr, xr))
t_br = timeit(lambda: ops.bernstein_forward_and_log_det_vjp(thr, xr, ybr, lbr))
bytes_f, bytes_b = (12 + 4 * order) * nr, (16 + 8 * order) * nr
print(f"rows n {nr}: forward {t_fr*1e3:.1f} us ({bytes_f/t_fr/1e6:.0f} GB/s), vjp {t_br*1e3:.1f} us ({bytes_b/t_br/1e6:.0f} GB/s)")
