"""evaluate_actions (the PPO update's differentiable masked log-prob + entropy) and ChoiceStore against plain float32
torch and against the reference's per-env loop (a2c_ppo_acktr/model.py:156-198, algo/ppo.py:51-66,
storage.py:19,48-57,133-145).  Floating point: tolerance 1e-5 (absolute, float32), as north_star states for the
sampler."""
import numpy as np
import pytest
import torch

from gymothelloenv_b200 import ChoiceStore, VecOthelloEnv, _cabi, evaluate_actions

pytestmark = pytest.mark.gpu
TOL = 1e-5


def _random_case(B, S, seed, dev):
    g = torch.Generator(device="cpu").manual_seed(seed)
    N, W = S * S, _cabi.words_per_board(S)
    logits = (torch.randn((B, N), generator=g) * 2.0).to(dev)
    dense = torch.rand((B, N), generator=g) < 0.25
    dense[::17] = False                                   # rows without any choice
    dense[1::17, :] = False
    dense[1::17, 5] = True                                # rows with exactly one choice
    legal = torch.zeros((W, B), dtype=torch.int64)
    for a in range(N):
        legal[a >> 6] |= dense[:, a].to(torch.int64) << (a & 63)
    # actions: mostly a legal one, sometimes an illegal / out-of-range one
    acts = torch.zeros(B, dtype=torch.int64)
    for i in range(B):
        ids = torch.nonzero(dense[i]).flatten()
        if len(ids) and i % 11 != 3:
            acts[i] = ids[int(torch.randint(len(ids), (1,), generator=g))]
        else:
            acts[i] = int(torch.randint(N, (1,), generator=g))
    acts[7] = N + 3
    return logits, dense.to(dev), legal.to(dev), acts.to(dev)


def _torch_reference(logits, dense, acts):
    """float32 torch restatement: log_softmax over the legal subset, 0 where the action is no legal choice; entropy
    of the full distribution (torch.distributions.Categorical, what FixedCategorical.entropy() is)."""
    N = logits.shape[1]
    masked = logits.masked_fill(~dense, float("-inf"))
    lsm = torch.log_softmax(masked, dim=1)
    in_range = (acts >= 0) & (acts < N)
    a = acts.clamp(0, N - 1)
    chosen_legal = in_range & dense.gather(1, a.unsqueeze(1)).squeeze(1)
    picked = lsm.gather(1, a.unsqueeze(1)).squeeze(1)
    logp = torch.where(chosen_legal, picked, torch.zeros_like(picked))
    ent = torch.distributions.Categorical(logits=logits).entropy().mean()
    return logp.unsqueeze(1), ent


@pytest.mark.parametrize("B,S", [(65536, 8), (4096, 6), (2048, 10), (1024, 16), (777, 5)])
def test_forward_and_backward_match_torch(B, S):
    dev = torch.device("cuda", 0)
    logits, dense, legal, acts = _random_case(B, S, 100 + S, dev)
    x1 = logits.clone().requires_grad_(True)
    x2 = logits.clone().requires_grad_(True)
    lp1, ent1 = evaluate_actions(x1, legal, acts, S)
    lp2, ent2 = _torch_reference(x2, dense, acts)
    assert lp1.shape == (B, 1)
    assert torch.allclose(lp1, lp2, atol=TOL, rtol=0)
    assert abs(float(ent1) - float(ent2)) < TOL
    g = torch.Generator(device="cpu").manual_seed(1)
    w = torch.randn((B, 1), generator=g).to(dev)
    (lp1 * w).sum().backward()
    (lp2 * w).sum().backward()
    assert torch.allclose(x1.grad, x2.grad, atol=TOL, rtol=0)
    no_grad_rows = (lp2.detach().squeeze(1) == 0) & ~dense.gather(1, acts.clamp(0, S * S - 1).unsqueeze(1)).squeeze(1)
    assert no_grad_rows.any() and float(x1.grad[no_grad_rows].abs().max()) == 0.0     # constant 0 -> all-zero row
    x1.grad = None
    x2.grad = None
    lp1, ent1 = evaluate_actions(x1, legal, acts, S)
    lp2, ent2 = _torch_reference(x2, dense, acts)
    (ent1 * 3.0).backward()
    (ent2 * 3.0).backward()
    assert torch.allclose(x1.grad, x2.grad, atol=TOL, rtol=0)


def test_clipped_ppo_surrogate_equals_the_per_env_loop():
    """algo/ppo.py:51-66 with the reference's evaluate_actions loop (model.py:163-172) on the host vs the batched path:
    same loss, same gradient of the logits.  The choices come from real positions through ChoiceStore."""
    dev = torch.device("cuda", 0)
    S, Bn, T = 8, 96, 6
    env = VecOthelloEnv(Bn, board_size=S, seed=4, auto_reset=True, obs_layout="planes4", obs_dtype=torch.float32)
    env.reset()
    store = ChoiceStore(T, Bn, S, dev)
    actions = torch.zeros((T, Bn, 1), dtype=torch.int64, device=dev)
    for t in range(T):
        store.insert(env._legal)
        env.step(policy="random")
        actions[t, :, 0] = env.actions_taken
    actions[2, 5, 0] = 0                                    # (0,0) is never legal on move 3: a "not in choices" row
    g = torch.Generator(device="cpu").manual_seed(3)
    idx = torch.randperm(T * Bn, generator=g)[:256]
    idx[0] = 2 * Bn + 5                                     # the 'action not in choices' sample is in the minibatch
    lists = store.as_lists(idx)
    for j, i in enumerate(idx.tolist()):                    # gather order = storage.py's flattening (t * B + b)
        t, b = divmod(i, Bn)
        words = store.data[t, :, b].cpu().tolist()
        assert lists[j] == [a for a in range(S * S) if (words[a >> 6] >> (a & 63)) & 1]
    logits = torch.randn((256, S * S), generator=g).to(dev)
    old_lp = (torch.randn((256, 1), generator=g) * 0.3 - 2.0).to(dev)
    adv = torch.randn((256, 1), generator=g).to(dev)
    acts_b = actions.view(-1, 1)[idx.to(dev)]
    clip, ent_coef = 0.2, 0.01

    def loss_of(lp, ent):
        ratio = torch.exp(lp - old_lp)
        surr1, surr2 = ratio * adv, torch.clamp(ratio, 1.0 - clip, 1.0 + clip) * adv
        return -torch.min(surr1, surr2).mean() - ent * ent_coef

    x1 = logits.clone().requires_grad_(True)
    lp1, ent1 = evaluate_actions(x1, store.gather(idx), acts_b, S)
    l1 = loss_of(lp1, ent1)
    l1.backward()
    # the reference's loop, on the host
    x2 = logits.cpu().clone().requires_grad_(True)
    rows = []
    for i in range(256):
        ch, a = lists[i], int(acts_b[i, 0])
        if len(ch) == 0 or a not in ch:
            rows.append(torch.tensor([[0.0]]))
        else:
            d = torch.distributions.Categorical(logits=x2[i][ch])
            rows.append(d.log_prob(torch.tensor([ch.index(a)])).view(1, 1))
    lp2 = torch.cat(rows)
    ent2 = torch.distributions.Categorical(logits=x2).entropy().mean()
    ratio = torch.exp(lp2 - old_lp.cpu())
    l2 = -torch.min(ratio * adv.cpu(), torch.clamp(ratio, 1.0 - clip, 1.0 + clip) * adv.cpu()).mean() - ent2 * ent_coef
    l2.backward()
    assert abs(float(l1) - float(l2)) < TOL
    assert torch.allclose(x1.grad.cpu(), x2.grad, atol=TOL, rtol=0)
    assert float(lp1[(lp2 == 0).squeeze(1).to(dev)].abs().sum()) == 0.0 and int((lp2 == 0).sum()) >= 1
