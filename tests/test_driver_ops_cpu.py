"""CPU: oracle restatements of apply_kl_penalty / compute_data_metrics against golden vectors produced by the
reference's own function bodies (oracle/make_golden.py::gen_driver)."""

import json

import numpy as np
import pytest
import torch

from conftest import load_golden
from oracle import port


def T(a):
    return torch.from_numpy(np.asarray(a))


@pytest.mark.parametrize("kind", ["kl", "abs", "mse", "low_var_kl"])
def test_kl_reward_golden(kind):
    g = load_golden("driver")
    rewards, cur = port.kl_reward(T(g["scores"]), T(g["old"]), T(g["ref"]), T(g["rmask"]), kind, 0.05)
    torch.testing.assert_close(rewards, T(g[f"klr|{kind}|rewards"]), rtol=1e-6, atol=1e-7)
    assert abs(cur - float(g[f"klr|{kind}|scalars"][0])) <= 1e-6 * max(1.0, abs(cur))


@pytest.mark.parametrize("critic", [1, 0])
def test_data_metrics_golden(critic):
    g = load_golden("driver")
    want = json.loads(str(g[f"dm|critic{critic}"]))
    rewards = T(g["klr|kl|rewards"])
    got = port.data_metrics(T(g["scores"]), rewards, T(g["adv"]), T(g["returns"]), T(g["values"]) if critic else None,
                            T(g["rmask"]), T(g["attn"]))
    assert set(got) == set(want)
    for k, v in want.items():
        assert abs(got[k] - v) <= 1e-5 * max(1.0, abs(v)), k
