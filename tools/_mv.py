import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, bench
from plutracer_b200 import GpuContext
for name in ("glass0", "minimal0", "refrac0"):
    c = bench.CONFIGS[name]; sc = bench.load_scene(name)
    ctx = GpuContext(0); ctx.upload(sc)
    w, h, n = c["w"], c["h"], c["n"]
    fb = torch.zeros(h, w, 3, dtype=torch.float32, device="cuda")
    for mv in (8, 5, 3, 2, 1):
        p = GpuContext.params(w, h, n, seed=bench.SEED, max_vertices=mv)
        best = 1e9
        for _ in range(6):
            st = ctx.render_device(p, fb.data_ptr(), torch.cuda.current_stream().cuda_stream)
            best = min(best, st.ms_device)
        print(name, "max_vertices", mv, "ms", round(best, 3), "waves", st.waves, "launches", st.kernel_launches, "rays", st.rays_extend + st.rays_shadow)
    ctx.close()
