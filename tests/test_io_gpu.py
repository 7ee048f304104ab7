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


def test_bad_checkpoints_leave_the_state_untouched(b200, tmp_path, monkeypatch):
    """Everything wrong with a checkpoint file is found before the first byte reaches the state: a truncated file, a ψ count
    that does not match the handle's PML and a ψ blob saved with another row pitch all raise and change nothing."""
    from fdtd_b200 import io as fio

    size = (100, 12, 10)                                   # rows padded to 128 cells
    w = _random_workload(b200, size, seed=4)
    w.pml = b200.PmlProfiles.graded(w.solver_config(), 3)

    def fresh(with_pml=True):
        ww = _random_workload(b200, size, seed=4)
        ww.pml = w.pml if with_pml else None
        return b200.FdtdB200Backend(exact_arithmetic=True).create_instance(ww.solver_config(), ww.domain()).create_state()

    a = fresh()
    a.set_source_table(w.source_points(), w.source_table(5))
    a.step(5)
    path = str(tmp_path / "ck.bin")
    fio.save_checkpoint(a, path)
    blob = open(path, "rb").read()

    b = fresh()
    b.set_source_table(w.source_points(), w.source_table(2))
    b.step(2)
    before = (b.read_field(b200.FieldComponent.E).copy(), b.read_field(b200.FieldComponent.H).copy(), b.tick(), b.time())

    def unchanged():
        return (np.array_equal(b.read_field(b200.FieldComponent.E), before[0]) and np.array_equal(b.read_field(b200.FieldComponent.H), before[1])
                and b.tick() == before[2] and b.time() == before[3])

    cut = str(tmp_path / "cut.bin")
    open(cut, "wb").write(blob[: len(blob) - 100])
    with pytest.raises(ValueError, match="truncated"):
        fio.load_checkpoint(b, cut)
    assert unchanged()
    nopml = fresh(with_pml=False)
    with pytest.raises(ValueError, match="ψ values"):
        fio.load_checkpoint(nopml, path)
    assert nopml.tick() == 0 and not np.any(nopml.read_field(b200.FieldComponent.E))
    monkeypatch.setenv("FDTD_B200_NO_ROW_PADDING", "1")
    c = fresh()
    assert c.row_pitch == 100 and a.row_pitch == 128
    with pytest.raises(ValueError, match="ψ values|rows of 128 cells"):      # the ψ arrays of a padded handle are larger, which is caught first
        fio.load_checkpoint(c, path)
    assert c.tick() == 0 and not np.any(c.read_field(b200.FieldComponent.H))
    monkeypatch.delenv("FDTD_B200_NO_ROW_PADDING")
    fio.load_checkpoint(b, path)                           # and the good file still loads
    assert b.tick() == 5 and np.array_equal(b.read_field(b200.FieldComponent.E), a.read_field(b200.FieldComponent.E))
