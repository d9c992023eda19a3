"""Design probe: zero-padded 3-D Toeplitz FFT pair, full cuFFT plan vs pruned (2-D transforms on the n nonzero
x-planes of an x-outermost layout, then one strided 1-D transform along x)."""
import torch, sys
B, n, F = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
dev = "cuda"
u = torch.zeros(B, F, F, F, dtype=torch.complex128, device=dev)
u[:, :n, :n, :n] = torch.randn(B, n, n, n, dtype=torch.complex128, device=dev)
t = torch.zeros(F, B, F, F, dtype=torch.complex128, device=dev)
t[:n, :, :n, :n] = u[:, :n, :n, :n].permute(1, 0, 2, 3)
def timeit(fn, reps=5):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): r = fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps, r
ms_full, rf = timeit(lambda: torch.fft.fftn(u, dim=(1, 2, 3)))
def pruned():
    a = torch.fft.fftn(t[:n], dim=(2, 3))          # (n, B, F, F): contiguous leading planes
    w = torch.zeros(F, B, F, F, dtype=torch.complex128, device=dev)
    w[:n] = a
    return torch.fft.fft(w, dim=0)
ms_pr, rp = timeit(pruned)
def pruned_parts():
    a = torch.fft.fftn(t[:n], dim=(2, 3))
    return a
ms_p1, _ = timeit(pruned_parts)
w = torch.zeros(F, B, F, F, dtype=torch.complex128, device=dev)
ms_p2, _ = timeit(lambda: torch.fft.fft(w, dim=0))
err = (rp.permute(1, 0, 2, 3) - rf).abs().max().item() / rf.abs().max().item()
print(f"B={B} n={n} F={F}: full 3-D {ms_full:.3f} ms | pruned total (with torch copies) {ms_pr:.3f} ms | 2-D on n planes {ms_p1:.3f} + strided 1-D along x {ms_p2:.3f} = {ms_p1+ms_p2:.3f} ms | rel err {err:.2e}")
