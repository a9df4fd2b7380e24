"""Scene ingest (csrc/ingest.cpp): heightmaps and Wavefront OBJ -> the reference's Object
array, plus the `factor` present pass on the GPU."""
import numpy as np
import pytest

import raytracing_cuda_b200 as rt

CUBE = """# unit cube, counter-clockwise seen from outside
v 0 0 0
v 1 0 0
v 1 1 0
v 0 1 0
v 0 0 1
v 1 0 1
v 1 1 1
v 0 1 1
f 1 4 3 2
f 5 6 7 8
f 1/1 2/2 6/3 5/4
f 2//1 3//1 7//1 6//1
f 3/1/1 4/2/1 8/3/1 7/4/1
f -4 -1 -5 -8
"""


def _normals(objs):
    a, b, c = (objs[k].astype(np.float64) for k in ("pos", "size", "third"))
    return np.cross(b - a, c - a), (a + b + c) / 3


def test_heightmap_cells_winding_and_material():
    hm = np.array([[0, 1, 0], [1, 3, 1], [0, 1, 0], [2, 2, 2]], dtype=np.float32)      # nz = 4, nx = 3
    objs = rt.heightmap_objects(hm, -1.0, 1.0, 10.0, 16.0)
    assert len(objs) == 2 * 2 * 3
    assert (objs["type"] == 2).all() and (objs["light"] == 0).all()
    n, _ = _normals(objs)
    assert (n[:, 1] > 0).all()                    # every triangle faces up, like the reference's island top
    # the first cell: (p0, p2, p1), (p0, p3, p2) with p0 = (x0, h00, z0)
    o = objs[0]
    assert tuple(o["pos"]) == (-1.0, 0.0, 10.0) and tuple(o["size"]) == (0.0, 3.0, 12.0) and tuple(o["third"]) == (0.0, 1.0, 10.0)
    o = objs[1]
    assert tuple(o["pos"]) == (-1.0, 0.0, 10.0) and tuple(o["size"]) == (-1.0, 1.0, 12.0) and tuple(o["third"]) == (0.0, 3.0, 12.0)
    # material template
    s = rt.Scene()
    tmpl = s.objects[0].copy()                    # the sea plane: mirror 0.6
    objs2 = rt.heightmap_objects(hm, -1.0, 1.0, 10.0, 16.0, material=tmpl)
    assert np.allclose(objs2["mirror"], 0.6) and (objs2["type"] == 2).all()
    with pytest.raises(ValueError):
        rt.heightmap_objects(np.zeros((1, 5), dtype=np.float32), 0, 1, 0, 1)


def test_obj_parse_cube():
    objs = rt.load_obj(CUBE)
    assert len(objs) == 12                        # six quads, fan-triangulated
    n, centroid = _normals(objs)
    outward = centroid - 0.5
    assert ((n * outward).sum(-1) > 0).all()      # winding kept: every face looks outward
    with pytest.raises(ValueError):
        rt.load_obj("v 0 0 0\nv 1 0 0\nf 1 2 9\n")
    assert len(rt.load_obj("# nothing\n")) == 0


@pytest.mark.gpu
def test_obj_scene_renders_and_bvh_equals_bruteforce():
    import gpu_util as gu
    from raytracing_cuda_b200 import _lib
    s = rt.Scene()
    cube = rt.load_obj(CUBE)
    for k in ("pos", "size", "third"):            # scale and move the cube onto the island
        cube[k] = cube[k] * 6.0 + np.array([-10.0, 0.0, 5.0], dtype=np.float32)
    objs = np.concatenate([s.objects, cube])
    r = gu.renderer()
    f = s.configure(view=5, hours=14.0, imgw=640, imgh=360, alias=True)
    img, ids = gu.render(r, s, f, objects=objs)
    assert (ids >= len(s.objects)).sum() > 500    # the cube is in view
    r.set_option(_lib.RT_OPT_TRAVERSAL, 1)
    img_b, ids_b = gu.render(r, s, f, objects=objs)
    r.set_option(_lib.RT_OPT_TRAVERSAL, 0)
    assert (ids == ids_b).all() and (img == img_b).all()


@pytest.mark.gpu
@pytest.mark.parametrize("factor", [1, 2, 3])
def test_present_factor_nearest_and_box(factor):
    import torch
    import gpu_util as gu
    r = gu.renderer()
    w, h = 96, 54
    rng = np.random.default_rng(factor)
    src = rng.integers(0, 1 << 24, size=(h * factor, w * factor), dtype=np.uint32)
    d_in = gu.to_dev(src)
    d_out = torch.zeros(w * h, dtype=torch.int32, device="cuda")
    r.present(d_in, w, h, factor, d_out)
    r.synchronize()
    got = gu.to_host_u32(d_out, (h, w))
    assert (got == src[factor // 2::factor, factor // 2::factor]).all()        # GL_NEAREST blit, main.cpp:317-318
    r.present(d_in, w, h, factor, d_out, box=True)
    r.synchronize()
    got = gu.to_host_u32(d_out, (h, w))
    ch = np.stack([(src >> s) & 255 for s in (0, 8, 16)], -1).astype(np.int64)
    blk = ch.reshape(h, factor, w, factor, 3).sum(axis=(1, 3))
    n = factor * factor
    ref = (blk + n // 2) // n
    assert (got == (ref[..., 0] | (ref[..., 1] << 8) | (ref[..., 2] << 16)).astype(np.uint32)).all()
