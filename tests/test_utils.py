"""tests/test_utils.py:7-26 (iou_loss known answers) + mesh / grid helpers."""
import numpy as np
import torch

from drtvam_b200.utils import iou_loss, load_ply, mesh_inside, reshape_grid, write_exr, write_ply


def test_iou():
    target = torch.tensor([1., 1., 0., 0.]).reshape(2, 2)
    assert iou_loss(torch.ones(2, 2), target) == 0.5
    assert iou_loss(torch.zeros(2, 2), target) == 0.
    assert iou_loss(torch.tensor([1., 0, 0, 0]).reshape(2, 2), target) == 0.5
    assert torch.isclose(iou_loss(torch.tensor([1., 1, 1, 0]).reshape(2, 2), target), torch.tensor(2 / 3))
    assert iou_loss(torch.tensor([0.6, 0.6, 0.6, 0]).reshape(2, 2), target) == 0.
    assert torch.isclose(iou_loss(torch.tensor([0.6, 0.6, 0.6, 0]).reshape(2, 2), target, threshold=0.5),
                         torch.tensor(2 / 3))


def test_reshape_grid():
    a = np.arange(5 * 2 * 3, dtype=np.float64).reshape(5, 2, 3)
    g = reshape_grid(a)
    assert g.shape == (6, 9, 1)
    assert np.array_equal(g[:2, :3, 0], a[0]) and np.array_equal(g[2:4, 3:6, 0], a[4])


def test_ply_roundtrip_and_inside(tmp_path):
    from drtvam_b200.synthetic import box_hole_mesh
    v, t = box_hole_mesh()
    path = str(tmp_path / "m.ply")
    write_ply(path, v, t)
    v2, t2 = load_ply(path)
    assert np.allclose(v, v2, atol=1e-5) and np.array_equal(t, t2)
    pts = torch.tensor([[30., 30., 0.], [-10., 0., 0.], [39., -39., 19.], [0., 0., 25.], [45., 0., 0.]])
    g = torch.Generator().manual_seed(0)
    d = torch.nn.functional.normalize(torch.randn(5, 3, generator=g), dim=-1)
    assert mesh_inside(pts, d, v, t).tolist() == [True, False, True, False, False]


def test_exr_writer(tmp_path):
    path = str(tmp_path / "a.exr")
    write_exr(path, np.random.rand(5, 7).astype(np.float32))
    raw = open(path, "rb").read()
    assert raw[:4] == bytes([0x76, 0x2f, 0x31, 0x01]) and len(raw) > 5 * 7 * 4


def test_obj_meshes_load_like_ply(tmp_path):
    """load_mesh accepts the two mesh formats the reference's configs name (optimize.py:41 passes the
    extension on as the Mitsuba shape type): the same cube as PLY and as OBJ (quads, `i/t/n` entries, negative
    indices) gives the same triangles' geometry; anything else is refused."""
    import pytest
    from drtvam_b200.synthetic import cube_mesh
    from drtvam_b200.utils import load_mesh
    v, t = cube_mesh(2.0)
    ply = str(tmp_path / "cube.ply")
    write_ply(ply, v, t)
    obj = tmp_path / "cube.obj"
    quads = [(0, 1, 3, 2), (4, 6, 7, 5), (0, 4, 5, 1), (2, 3, 7, 6), (0, 2, 6, 4), (1, 5, 7, 3)]
    lines = ["# cube"] + [f"v {a} {b} {c}" for a, b, c in v] + ["vn 0 0 1", "vt 0 0"]
    for k, q in enumerate(quads):
        if k % 2 == 0:
            lines.append("f " + " ".join(f"{i + 1}/1/1" for i in q))
        else:
            lines.append("f " + " ".join(str(i - len(v)) for i in q))      # relative indices
    obj.write_text("\n".join(lines) + "\n")
    v1, t1 = load_mesh(ply)
    v2, t2 = load_mesh(str(obj))
    assert t2.shape == (12, 3) and np.allclose(v1, v2)

    def tri_set(vv, tt):
        return sorted(tuple(sorted(map(tuple, np.round(vv[tri], 9)))) for tri in tt)
    assert tri_set(v1, t1) == tri_set(v2, t2)
    with pytest.raises(ValueError):
        load_mesh(str(tmp_path / "cube.stl"))
