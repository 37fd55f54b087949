"""Tensor-format contract of the env adaptor (reference env_packer.py) on a synthetic vec-env."""
import numpy as np
import pytest
import torch

from microbeast_b200 import buffers as mbuf, synth
from microbeast_b200.env_packer import Env_Packer, SyntheticVecEnv, _format_action_masks, _format_obs
from oracle import refload


def test_formatters_match_reference_shapes():
    obs = np.random.RandomState(0).randint(0, 2, (3, 8, 8, 27)).astype(np.int32)
    mask = np.random.RandomState(1).randint(0, 2, (3, 78 * 64)).astype(np.uint8)
    o, m = _format_obs(obs), _format_action_masks(mask)
    assert o.shape == (1, 1, 3, 8, 8, 27) and o.dtype == torch.float32        # env_packer.py:8-10
    assert m.shape == (1, 3, 78 * 64) and m.dtype == torch.uint8             # env_packer.py:12-14
    assert torch.equal(o[0, 0], torch.from_numpy(obs).float())
    assert _format_obs(obs, torch.uint8).dtype == torch.uint8


def test_int32_masks_of_gym_microrts_become_uint8():
    """gym-microRTS ``get_action_mask()`` returns int32 and the reference only casts inside ``CategoricalMasked``
    (``model.py:40``); the packer normalises once so the rollout store and the kernels see one byte per entry."""
    rs = np.random.RandomState(3)
    for dtype in (np.int32, np.int64, np.bool_, np.float32, np.uint8):
        raw = (rs.randint(0, 3, (2, 78 * 16)) * (rs.rand(2, 78 * 16) < 0.5)).astype(dtype)     # values 0 / 1 / 2
        m = _format_action_masks(raw)
        assert m.dtype == torch.uint8 and m.shape == (1, 2, 78 * 16)
        assert torch.equal(m[0] != 0, torch.from_numpy(raw != 0))

    class Int32Env(SyntheticVecEnv):
        def get_action_mask(self):
            return super().get_action_mask().astype(np.int32)

    packer = Env_Packer(Int32Env(2, 8, seed=5), 0, "")
    out = packer.initial()
    assert out["action_mask"].dtype == torch.uint8
    out = packer.step(torch.zeros((2, 7 * 64), dtype=torch.int64))
    assert out["action_mask"].dtype == torch.uint8


def test_packer_outputs_fit_the_rollout_buffers(tmp_path):
    n_envs, T = 3, 4
    env = SyntheticVecEnv(n_envs, 8, seed=2, episode_len=3)
    packer = Env_Packer(env, 0, str(tmp_path / "exp"))
    bufs = mbuf.create_buffers(2, n_envs, 1, T, (8, 8, 27))
    out = packer.initial()
    assert set(out) == {"obs", "reward", "done", "ep_return", "ep_step", "last_action", "action_mask"}
    for key in out:
        bufs[key][0][0, ...] = out[key]                                     # microbeast.py:73-75
    for t in range(T):
        action = synth.make_valid_actions(out["action_mask"][0], gen=torch.Generator().manual_seed(t))
        out = packer.step(action)
        assert out["obs"].shape == (1, 1, n_envs, 8, 8, 27) and out["reward"].shape == (1, n_envs)
        assert out["done"].shape == (1, n_envs) and out["action_mask"].shape == (1, n_envs, 78 * 64)
        for key in out:
            bufs[key][0][t + 1, ...] = out[key]                             # microbeast.py:95-96
    assert torch.equal(bufs["obs"][0].sum(-1), torch.full((T + 1, n_envs, 8, 8), 5, dtype=torch.uint8))
    rows = (tmp_path / "exp.csv").read_text().strip().splitlines()
    assert rows and all(len(r.split(",")) == 3 for r in rows)               # env_packer.py:69-75
    # actions drawn from the masks the env handed out are all accepted
    assert env.invalid_actions == 0


@pytest.mark.skipif(not refload.available(), reason="reference checkout not present")
def test_formatters_equal_the_reference_functions():
    import importlib, sys
    refload.load()
    ref = importlib.import_module("env_packer")
    obs = np.random.RandomState(0).randint(0, 2, (2, 8, 8, 27)).astype(np.int32)
    mask = np.random.RandomState(1).randint(0, 2, (2, 78 * 64)).astype(np.uint8)
    assert torch.equal(ref._format_obs(obs), _format_obs(obs))
    assert torch.equal(ref._format_action_masks(mask), _format_action_masks(mask))
