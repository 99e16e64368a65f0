"""CPU: host-side workspace logic of the kernel schedules (no device needed) — the row-range views a stack gets when the batch is
cut into sample-range chains (transformer.WorkspacePart, iq.DEC_SPLIT)."""
import torch


def test_workspace_parts_are_row_ranges_of_the_parent_buffers():
    from blt_vqg_b200 import transformer as tf
    ws = tf.Workspace(torch.device("cpu"))
    parts = [tf.WorkspacePart(ws, i, 2) for i in range(2)]
    a0, a1 = parts[0].buf("x", 6, 4), parts[1].buf("x", 6, 4)
    full = ws["x"]
    assert tuple(full.shape) == (12, 4)
    a0.fill_(1.0)
    a1.fill_(2.0)
    assert torch.equal(full[:6], torch.ones(6, 4)) and torch.equal(full[6:], torch.full((6, 4), 2.0))
    # the same request again returns the same storage (addresses stay stable: CUDA-graph replays)
    assert parts[0].buf("x", 6, 4).data_ptr() == a0.data_ptr() and parts[1]["x"].data_ptr() == a1.data_ptr()
    # other dtypes / ranks: the leading dimension carries the batch
    ids = parts[1].buf("ids", 3, 5, dtype=torch.int32)
    assert tuple(ws["ids"].shape) == (6, 5) and ids.data_ptr() == ws["ids"][3:].data_ptr()


def test_workspace_part_meta_is_private_except_operand_widths():
    from blt_vqg_b200 import transformer as tf
    ws = tf.Workspace(torch.device("cpu"))
    p0, p1 = tf.WorkspacePart(ws, 0, 2), tf.WorkspacePart(ws, 1, 2)
    p0.meta["dec.in"], p1.meta["dec.in"] = "a", "b"
    assert p0.meta["dec.in"] == "a" and p1.meta["dec.in"] == "b" and "dec.in" not in ws.meta
    p0.meta[("actK", "dec.lnf.y")] = 512   # describes the shared full-size planes: whoever runs after the join reads the parent
    assert ws.meta[("actK", "dec.lnf.y")] == 512
    # planes made by the parts are row ranges of the parent's planes
    hi0 = p0.buf("dec.lnf.y.hi", 4, 512, dtype=torch.float16)
    p1.buf("dec.lnf.y.hi", 4, 512, dtype=torch.float16)
    full = ws.act("dec.lnf.y")
    assert full.f32 is None and tuple(full.hi.shape) == (8, 512) and full.lo is None
    assert p0.act("dec.lnf.y").hi.data_ptr() == hi0.data_ptr()
