"""Minimal GPU probe of b200vr_params::pt_child_instance_xfm (no torch, no oracle): mean radiance of a path-traced 40^3 noise
volume with the switch off / on. Oracle values at 32 accumulated frames, 96x64: 0.356 (off) / 0.255 (on)."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
t0 = time.time()
import ovr_b200  # noqa: E402

S = ovr_b200.scenes
out = []
for flag in (0, 1):
    sc = S.noise_scene(40, tf="pt", camera="front")
    r = ovr_b200.create_renderer("b200")
    r.pt_child_instance_xfm = bool(flag)
    r.set_fbsize((96, 64))
    r.set_path_tracing(True)
    r.set_frame_accumulation(True)
    r.init(sc, sc.camera)
    r.commit()
    for _ in range(32):
        r.render()
    img = r.download()[0]
    hit = img[..., 3] > 0
    out.append(f"pt_child_instance_xfm={flag}: mean radiance {float(img[hit][:, :3].mean()):.4f} over {int(hit.sum())} pixels, "
               f"{r.last_stats['delta_samples']} collisions in the last frame, {time.time() - t0:.1f} s")
    r.close()
print("\n".join(out))
