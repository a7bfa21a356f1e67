# host-side costs around a short render: numpy fill, pageable vs pinned cudaMemcpy of one float image (glass0 1280x960, room 1920x1080)
import time, numpy as np, torch
for (w, h) in ((512, 512), (1280, 960), (1920, 1080)):
    n = w * h * 3
    a = np.zeros(n, np.float32); a.fill(0)
    p = torch.zeros(n, dtype=torch.float32).pin_memory()
    d = torch.zeros(n, dtype=torch.float32, device="cuda")
    ta = torch.from_numpy(a)
    def t(fn, reps=20):
        fn(); torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(reps): fn()
        torch.cuda.synchronize()
        return (time.perf_counter() - t0) / reps * 1e3
    print(f"{w}x{h} {n*4/1e6:.1f} MB: fill {t(lambda: a.fill(0)):.3f} ms, pageable h2d {t(lambda: d.copy_(ta)):.3f}, pageable d2h {t(lambda: ta.copy_(d)):.3f}, "
          f"pinned h2d {t(lambda: d.copy_(p, non_blocking=True)):.3f}, pinned d2h {t(lambda: p.copy_(d, non_blocking=True)):.3f}, device memset {t(lambda: d.zero_()):.3f}")
