"""Pins oracle/train_oracle.py against the third-party code the reference calls:
torch.optim.Adam on the CPU (GaussianPointTrainer.py:131-134,183-184) and the torch statements of
GaussianPointAdaptiveController.update (AC.py:130-146)."""
import numpy as np
import torch

from oracle import train_oracle as TO


def test_adam_matches_torch_optim():
    rng = np.random.default_rng(0)
    p0 = rng.normal(0, 1, (257, 5))
    for dtype, tol in ((torch.float64, 1e-13), (torch.float32, 2e-6)):
        p = torch.nn.Parameter(torch.from_numpy(p0).to(dtype))
        opt = torch.optim.Adam([p], lr=3e-3, betas=(0.9, 0.999), eps=1e-8)
        np_dtype = np.float64 if dtype == torch.float64 else np.float32
        po, m, v = p0.astype(np_dtype), np.zeros_like(p0, np_dtype), np.zeros_like(p0, np_dtype)
        for step in range(1, 8):
            g = rng.normal(0, 1, p0.shape) * (0.0 if step == 4 else 1.0)  # one all-zero gradient
            p.grad = torch.from_numpy(g).to(dtype)
            opt.step()
            po, m, v = TO.adam_step(po, g.astype(np_dtype), m, v, 3e-3, 0.9, 0.999, 1e-8, step, dtype=np_dtype)
            assert np.abs(p.detach().numpy() - po).max() <= tol * max(1.0, np.abs(po).max())


def test_controller_update_matches_torch_statements():
    rng = np.random.default_rng(1)
    n, m = 500, 180
    acc = TO.new_accumulators(n)
    t_acc = {k: torch.from_numpy(v.copy()) for k, v in acc.items()}
    for it in range(3):
        ids = np.sort(rng.choice(n, m, replace=False)).astype(np.int32)
        px = rng.integers(0, 50, m).astype(np.int32)
        mag = rng.uniform(0, 1, m).astype(np.float32)
        mag[px == 0] = 0.0                      # 0/0: the NaN the reference zeroes
        gx = rng.normal(0, 1, (m, 3)).astype(np.float32)
        TO.controller_update(acc, ids, px, mag, gx)
        # AC.py:130-146, verbatim semantics
        tid, tpx, tmag, tgx = (torch.from_numpy(a) for a in (ids.astype(np.int64), px, mag, gx))
        t_acc["accumulated_num_in_camera"][tid] += 1
        t_acc["accumulated_num_pixels"][tid] += tpx
        t_acc["accumulated_view_space_position_gradients"][tid] += tmag
        avg = tmag / tpx
        avg[torch.isnan(avg)] = 0
        t_acc["accumulated_view_space_position_gradients_avg"][tid] += avg
        t_acc["accumulated_position_gradients"][tid] += tgx
        t_acc["accumulated_position_gradients_norm"][tid] += tgx.norm(dim=1)
    for k in acc:
        a, b = acc[k], t_acc[k].numpy()
        if a.dtype == np.int32:
            assert np.array_equal(a, b), k
        else:
            assert np.allclose(a, b, rtol=1e-6, atol=1e-7), k
