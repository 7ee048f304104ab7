"""Projection passes into the GIF target and checkpoint save / load / resume on the device."""
import numpy as np
import pytest

from helpers import random_fields
from test_parity_gpu import _random_workload

pytestmark = pytest.mark.gpu


def test_projection_pass_into_gif_target(b200, tmp_path):
    from fdtd_b200 import io as fio

    w = _random_workload(b200, (32, 24, 8), seed=5)
    inst = b200.FdtdB200Backend().create_instance(w.solver_config(), w.domain())
    st = inst.create_state()
    e0, h0 = random_fields(w, seed=2)
    st.write_field(b200.FieldComponent.E, e0)
    st.write_field(b200.FieldComponent.H, h0)
    st.set_source_table(w.source_points(), w.source_table(12))
    cmap = np.eye(4, dtype=np.float32) * 0.3
    cmap[:3, 3], cmap[3, 3] = 0.5, 1.0
    params = b200.ProjectionParameters(np.eye(4, dtype=np.float32), b200.FieldComponent.E, cmap)
    path = str(tmp_path / "e.gif")
    direct = []
    with fio.GifEncoder(path, (32, 24)) as gif:
        to_gif = inst.create_projection(st, gif, params)
        to_img = inst.create_projection(st, b200.ImageTarget(32, 24), params)
        for _ in range(3):
            st.step(4)
            p = inst.begin_projection_pass(st)
            p.add_projection(to_gif)
            p.add_projection(to_img)
            p.finish()
            direct.append(to_img.target.image.copy())
    frames = fio.decode_gif(path)
    assert len(frames) == 3
    pal = fio._cube_palette()
    for f, img in zip(frames, direct):
        assert np.array_equal(f, pal[fio._quantise(img)])     # the GIF holds exactly what the image target received
    assert all(len(np.unique(d[..., :3])) > 8 for d in direct) and not np.array_equal(direct[0], direct[2])


@pytest.mark.parametrize("fused", [False, True])
def test_checkpoint_resume_is_bit_exact(b200, tmp_path, fused):
    from fdtd_b200 import io as fio

    size = (128, 12, 10)
    w = _random_workload(b200, size, seed=9)
    tab = w.source_table(20)

    def fresh():
        st = b200.FdtdB200Backend(exact_arithmetic=True, fused_step=fused).create_instance(w.solver_config(), w.domain()).create_state()
        return st

    ref = fresh()
    ref.set_source_table(w.source_points(), tab)
    ref.step(20)
    a = fresh()
    a.set_source_table(w.source_points(), tab)
    a.step(7)                                                  # odd number of steps: the fused ping-pong parity differs on resume
    path = str(tmp_path / "ck.bin")
    fio.save_checkpoint(a, path, planes_per_chunk=3)
    b = fresh()
    fio.load_checkpoint(b, path, planes_per_chunk=4)
    assert b.tick() == 7 and b.time() == a.time()
    b.set_source_table(w.source_points(), tab[7:])
    b.step(13)
    assert b.tick() == 20 and b.time() == ref.time()
    for comp in (b200.FieldComponent.E, b200.FieldComponent.H):
        assert np.array_equal(b.read_field(comp), ref.read_field(comp))
    other = b200.FdtdB200Backend().create_instance(_random_workload(b200, (128, 12, 9), seed=9).solver_config(),
                                                   _random_workload(b200, (128, 12, 9), seed=9).domain()).create_state()
    with pytest.raises(ValueError):
        fio.load_checkpoint(other, path)
