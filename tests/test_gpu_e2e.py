"""-m gpu: the path end to end -- train steps reduce the loss, the render loop matches with either compaction,
and the reference's own caller code (legacy/nerf/renderer.py semantics restated in trainer.py) runs on the operators."""
import numpy as np
import pytest
import torch

import cases

pytestmark = pytest.mark.gpu


def _trainer(bound=1.0):
    from focnerf_b200.trainer import Trainer
    scene, cascade, grid, bits, thresh = cases.scene_bitfield(bound)
    tr = Trainer(bound=bound, density_grid=torch.from_numpy(grid))
    assert np.array_equal(tr.density_bitfield.cpu().numpy(), bits)
    return scene, tr


def test_train_steps_reduce_loss():
    scene, tr = _trainer()
    losses = []
    for it in range(60):
        o, d = cases.camera_rays(4096, 100 + it)
        o, d = torch.from_numpy(o).cuda(), torch.from_numpy(d).cuda()
        gt = scene.render_gt(o, d)
        losses.append(float(tr.train_step(o, d, gt)))
        if (it + 1) % 16 == 0:
            tr.update_mean_count()
    assert all(np.isfinite(losses)), losses
    assert np.mean(losses[-10:]) < 0.5 * np.mean(losses[:5]), (losses[:5], losses[-10:])


def test_render_loop_compaction_equivalence():
    scene, tr = _trainer()
    o, d = cases.camera_rays(20000, 7)
    o, d = torch.from_numpy(o).cuda(), torch.from_numpy(d).cuda()
    a = tr.render(o, d, device_compaction=True)
    b = tr.render(o, d, device_compaction=False)
    for x, y in zip(a, b):
        assert torch.equal(torch.nan_to_num(x), torch.nan_to_num(y))
