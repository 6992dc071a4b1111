import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np
import common
from elevenrender_b200 import host
name = sys.argv[1] if len(sys.argv) > 1 else "grid"
spp = int(sys.argv[2]) if len(sys.argv) > 2 else 1
spec = common.make_spec(name)
def render(kernel):
    pars = host.RenderParameters(spec["x_res"], spec["y_res"], spp, rng_order=host.ER_RNG_RTL, kernel=kernel, spp_per_launch=spp)
    rm = host.RenderingManager(pars)
    rm.upload(common.to_ours(spec))
    t0 = time.time()
    rm.render_samples(spp)
    dt = time.time() - t0
    p = np.stack([rm.get_pass(q) for q in ("beauty", "normal")])
    rng, sa = rm.get_pixel_state()
    rm.close()
    return p, rng, sa, dt
a = render(0); b = render(3)
diff = (a[0].view(np.uint32) != b[0].view(np.uint32)).any(axis=(0, 3))
print(name, spp, "pixels differing", int(diff.sum()), "of", diff.size, "rng differ", int((a[1] != b[1]).sum()), "sa differ", int((a[2] != b[2]).sum()), "time k0 %.3f k3 %.3f" % (a[3], b[3]))
ys, xs = np.nonzero(diff)
print("first differing pixels (y,x):", list(zip(ys[:12].tolist(), xs[:12].tolist())))
if diff.any():
    y, x = ys[0], xs[0]
    print("k0", a[0][:, y, x], "k3", b[0][:, y, x])
