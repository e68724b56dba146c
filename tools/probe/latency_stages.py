"""Stage times of the single-frame call (batch 1, 752x480, nFeatures 1000): where the 0.16 ms go."""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path[:0] = [ROOT, os.path.join(ROOT, "orb-slam3-noted_b200")]
import orbfe
W, H = 752, 480
for B in (1, 2, 8, 64):
    ex = orbfe.ORBextractor(1000, 1.2, 8, 20, 7, max_width=W, max_height=H, max_batch=B)
    frames = np.stack([orbfe.synth.frame(1000 + i, W, H) for i in range(B)])
    for _ in range(5):
        ex.extract_batch(frames, (0, 1000))
    t0 = time.perf_counter()
    for _ in range(200):
        ex.extract_batch(frames, (0, 1000))
    wall = (time.perf_counter() - t0) / 200 * 1e3
    ex.set_profiling(True)
    for _ in range(50):
        ex.extract_batch(frames, (0, 1000))
    ms, calls = ex.stage_times()
    ex.set_profiling(False)
    print(f"batch {B}: wall {wall:.3f} ms per call; stages (ms per call, eager single stream):",
          {k: round(v / calls, 4) for k, v in ms.items()}, "sum", round(sum(ms.values()) / calls, 4))
