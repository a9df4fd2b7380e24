"""Second process of test_peer_frame_two_processes: opens the owner's peer frame (handle in argv) and renders
shard 1 of 2 into it for three frames."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import raytracing_cuda_b200 as rt  # noqa: E402
from raytracing_cuda_b200 import sharding  # noqa: E402

handle = bytes.fromhex(sys.argv[1])
w, h, tw, th, fused = (int(v) for v in sys.argv[2:7])
r = rt.Renderer(0)
r.set_sky(rt.synthetic_sky(256, 128))
s = rt.Scene()
f = s.configure(view=5, hours=18.0, imgw=w, imgh=h, alias=True)
r.set_scene(s.objects, s.lights)
pf = sharding.PeerFrame(r, 1, 2, w * h, lambda _: handle)
shard = rt.make_shard(2, 1, tw, th)
print("ready", flush=True)
for seq in (1, 2, 3):
    pf.render(f, shard, seq, fused=bool(fused))
r.synchronize()
err = pf.error()
pf.close()
sys.exit(3 if err else 0)
