"""Probe: cost of the 479-wide over-arch layer's three GEMMs with dense (pitch 479) vs padded (pitch 480) rows."""
import torch

torch.backends.cuda.matmul.allow_tf32 = True


def t(fn, n=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3


B, K, N = 65536, 479, 1024
for pitch in (479, 480):
    x = torch.randn(B, pitch, device="cuda")[:, :K]
    w = torch.randn(N, pitch, device="cuda")[:, :K]
    b = torch.randn(N, device="cuda")
    gy = torch.randn(B, N, device="cuda")
    gx = torch.empty(B, pitch, device="cuda")[:, :K]
    gw = torch.empty(N, pitch, device="cuda")[:, :K]
    print(f"pitch {pitch}: fwd {t(lambda: torch.addmm(b, x, w.t())):7.1f} us   "
          f"dX {t(lambda: torch.mm(gy, w, out=gx)):7.1f} us   dW {t(lambda: torch.mm(gy.t(), x, out=gw)):7.1f} us")
    y = torch.nn.functional.linear(x, w, b)
    print("   F.linear on the view:", f"{t(lambda: torch.nn.functional.linear(x, w, b)):7.1f} us", y.shape)
for kp in (13, 16):
    x = torch.randn(B, kp, device="cuda")[:, :13]
    w = torch.randn(512, kp, device="cuda")[:, :13]
    gy = torch.randn(B, 512, device="cuda")
    gw = torch.empty(512, kp, device="cuda")[:, :13]
    print(f"dense-in pitch {kp}: fwd {t(lambda: torch.mm(x, w.t())):7.1f} us  dW {t(lambda: torch.mm(gy.t(), x, out=gw)):7.1f} us")
