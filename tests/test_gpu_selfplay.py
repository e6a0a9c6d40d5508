"""GPU: the PPO self-play vec-env (gymothelloenv_b200.envs.PPOEnvs: oth_selfplay_* kernels)
and the legal-move-restricted sampler (oth_masked_categorical), through the C ABI.

Parity bar: every tensor PPOEnvs.step returns is bit-identical to what the LIVE reference
returned (tests/golden/selfplay.npz: envs.PPOEnvs + forked subproc_worker processes) under
the same deterministic stand-in policy, and to the CPU restatement on larger random batches.
The sampler is floating point: log-probabilities within 1e-5 (absolute) of a float32 torch
log_softmax over the legal subset; sampled actions must sit in the inverse-CDF interval the
independent numpy Philox draw selects (boundary slack 1e-5 of the total mass).
"""
import numpy as np
import pytest
import torch

from oracle.selfplay_oracle import SelfPlayOracle, rule_base_games
from tests import golden_io, philox_ref
from tests.selfplay_fake import FakePolicy

pytestmark = pytest.mark.gpu

CASES = golden_io.load_selfplay()


def _mk(B, S, disk, **kw):
    from gymothelloenv_b200 import PPOEnvs
    return PPOEnvs(B, None, FakePolicy(S), "cuda:0", board_size=S, num_disk_as_reward=disk, **kw)


def _unpack(ret):
    obs, actions, rewards, dones, infos, outputs, masks, bad_masks = ret
    return dict(obs=obs.cpu().numpy().astype(np.int8), actions=actions.cpu().numpy().reshape(-1).astype(np.int64),
                rewards=rewards.cpu().numpy().reshape(-1).astype(np.int32), dones=dones.cpu().numpy().astype(np.uint8),
                types=infos.type.cpu().numpy(), choices=infos.choices_mask.cpu().numpy().astype(np.uint8),
                outputs=outputs.cpu().numpy(), masks=masks.cpu().numpy().reshape(-1), bad_masks=bad_masks.cpu().numpy().reshape(-1))


@pytest.mark.parametrize("c", range(len(CASES)))
def test_ppoenvs_reproduces_live_reference(c):
    case = CASES[c]
    S, B = case["S"], case["B"]
    envs = _mk(B, S, case["disk"], initial_rand_steps=0)
    k = 0
    for g in range(case["games"]):
        envs.reset(protagonists=np.full(B, case["protagonist"][g]))
        while True:
            got = _unpack(envs.step(None))
            for key in ("obs", "actions", "rewards", "dones", "types", "choices", "outputs", "masks", "bad_masks"):
                assert np.array_equal(got[key], case[key][k]), (c, g, k, key)
            k += 1
            if got["dones"].all():
                break
    assert k == len(case["game"])
    envs.over()


@pytest.mark.parametrize("S,disk,legacy", [(8, True, True), (8, False, False), (6, True, True), (5, True, False), (10, True, True),
                                           (13, True, True), (16, False, True)])
def test_ppoenvs_vs_oracle_mixed_protagonists(S, disk, legacy):
    B = 384
    rng = np.random.RandomState(S * 7 + disk)
    envs = _mk(B, S, disk, legacy_stale_state=legacy)
    orc = SelfPlayOracle(B, S, disk_reward=disk, legacy_stale_state=legacy)
    pol = FakePolicy(S)
    for g in range(2):
        prot = rng.choice([-1, 1], size=B)
        envs.reset(protagonists=prot)
        orc.reset(prot)
        rounds = 0
        while True:
            got = _unpack(envs.step(None))
            want = orc.step(pol)
            for key in ("obs", "actions", "rewards", "dones", "types", "choices", "outputs"):
                assert np.array_equal(got[key], want[key]), (g, rounds, key)
            rounds += 1
            if got["dones"].all():
                break
        assert rounds > S * S // 2 - 4


def test_ppoenvs_opening_override_keeps_learner_action():
    """With an opening budget the env replaces the applied action by a uniform legal move
    (othello.py:70-73) but the tuple still carries the learner's own action (:289,306)."""
    S, B, seed, init = 8, 256, 77, 6
    envs = _mk(B, S, True, initial_rand_steps=init, seed=seed)
    orc = SelfPlayOracle(B, S, disk_reward=True)
    pol = FakePolicy(S)
    prot = np.where(np.arange(B) % 2 == 0, -1, 1)
    envs.reset(protagonists=prot)
    orc.reset(prot)
    cnt, mx = [x.cpu().numpy().astype(np.int64) for x in envs.vec.rand_steps()]
    assert (cnt == 0).all() and set(np.unique(mx)) <= {0, 2, 4, 6} and len(np.unique(mx)) > 1
    state = dict(cnt=cnt.copy(), ctr=envs.vec._ctr.cpu().numpy().astype(np.int64))
    gid = np.arange(B, dtype=np.int64)

    def hook(applied, need):
        _, legal, _, _, _, n_moves = orc.orc.export()
        ov = need & (state["cnt"] < mx)
        r = philox_ref.draw(seed, gid, 0, state["ctr"])
        idx = philox_ref.bounded(r, np.maximum(n_moves, 1))
        order = np.cumsum(legal, axis=1) - 1
        pick = ((order == idx[:, None]) & (legal == 1)).argmax(1)
        out = np.where(ov, pick, applied)
        state["cnt"] += ov
        state["ctr"] += ov
        return out

    overridden_rounds = 0
    while True:
        got = _unpack(envs.step(None))
        want = orc.step(pol, applied_actions_hook=hook)
        for key in ("obs", "actions", "rewards", "dones", "types", "choices", "outputs"):
            assert np.array_equal(got[key], want[key]), key
        overridden_rounds += 1
        if got["dones"].all():
            break
    assert np.array_equal(envs.vec.rand_steps()[0].cpu().numpy(), mx)


@pytest.mark.parametrize("c", [i for i, cs in enumerate(CASES) if cs["test_games"]])
def test_rule_base_game_matches_live_reference(c):
    case = CASES[c]
    S, B = case["S"], case["B"]
    per = case["test_games"] // B
    envs = _mk(B, S, case["disk"], test_rand_steps=0)
    prot = [int(p) for p in case["protagonist"][case["games"]:case["games"] + per]]
    assert envs.test("greedy", case["test_games"], None, protagonists=prot) == case["test"]


def test_rule_base_game_vs_oracle_and_random_opponent():
    S, B, per = 8, 200, 2
    rng = np.random.RandomState(5)
    prot = rng.choice([-1, 1], size=(per, B))
    envs = _mk(B, S, True, test_rand_steps=0)
    games, wins = envs.test("greedy", per * B, None, protagonists=prot)
    want = rule_base_games(FakePolicy(S), "greedy", per, S, prot, disk_reward=True)
    assert (games, wins) == (per * B, int(want.sum()))
    # one forward pass per LEARNER ply (envs.py:233-250 serves one request of every worker per act), not one per ply:
    # a game has at most (S*S - 4) plies, about half of them the learner's
    assert 0 < envs.test_policy_calls <= per * 0.6 * (S * S - 4)
    # 'rand' opponent + opening budget: statistical only (Philox vs numpy MT streams)
    envs = _mk(B, S, True, test_rand_steps=10)
    games, wins = envs.test("rand", 4 * B, None)
    assert games == 4 * B and 0.2 * games < wins < 0.8 * games


# ------------------------------------------------------------------------------ sampler
def _random_case(B, S, seed, empty_frac=0.05):
    from gymothelloenv_b200 import _cabi
    rng = np.random.RandomState(seed)
    N, W = S * S, _cabi.words_per_board(S)
    logits = (rng.randn(B, N) * 3).astype(np.float32)
    legal = rng.rand(B, N) < rng.uniform(0.02, 0.4, size=(B, 1))
    legal[rng.rand(B) < empty_frac] = False
    single = rng.rand(B) < 0.05
    legal[single] = False
    legal[np.nonzero(single)[0], rng.randint(0, N, size=int(single.sum()))] = True
    words = np.zeros((W, B), dtype=np.uint64)
    for a in range(N):
        words[a >> 6] |= legal[:, a].astype(np.uint64) << np.uint64(a & 63)
    return logits, legal, words.view(np.int64)


def _torch_logp(logits, legal):
    """float32 reference: FixedCategorical(logits=tensor[choices]).log_probs (model.py:69-85)"""
    x = torch.from_numpy(logits)
    m = torch.from_numpy(legal)
    return torch.log_softmax(x.masked_fill(~m, float("-inf")), dim=1).numpy()


@pytest.mark.parametrize("S", [8, 6, 10, 12, 5, 14, 16])
def test_masked_categorical_modes(S):
    from gymothelloenv_b200 import masked_categorical
    B, seed = 5000, 1234
    logits, legal, words = _random_case(B, S, S)
    lg, wd = torch.from_numpy(logits).cuda(), torch.from_numpy(words).cuda()
    ref = _torch_logp(logits, legal)
    has = legal.any(1)
    # evaluate: every legal action of a few envs, an illegal one, and the empty case
    acts = np.where(has, legal.argmax(1), 3)
    a, lp = masked_categorical(lg, wd, S, "evaluate", actions=torch.from_numpy(acts).cuda())
    lp = lp.cpu().numpy()
    assert np.array_equal(a.cpu().numpy(), acts)
    assert np.abs(lp[has] - ref[has, acts[has]]).max() < 1e-5
    assert (lp[~has] == 0).all()
    illegal = (~legal).argmax(1)
    ok_illegal = ~legal[np.arange(B), illegal]
    _, lp2 = masked_categorical(lg, wd, S, "evaluate", actions=torch.from_numpy(illegal).cuda())
    assert (lp2.cpu().numpy()[ok_illegal] == 0).all()               # model.py:172-173
    _, lp3 = masked_categorical(lg, wd, S, "evaluate", actions=torch.full((B,), -5, dtype=torch.int64).cuda())
    assert (lp3.cpu().numpy() == 0).all()
    # mode: first argmax over the legal subset
    a, lp = masked_categorical(lg, wd, S, "mode")
    a, lp = a.cpu().numpy(), lp.cpu().numpy()
    want = np.where(legal, logits, -np.inf).argmax(1)
    assert np.array_equal(a[has], want[has]) and (a[~has] == 0).all() and (lp[~has] == 0).all()
    assert np.abs(lp[has] - ref[has, want[has]]).max() < 1e-5
    # sample: inverse CDF in ascending action order driven by one Philox word
    base = 1 << 20
    a, lp = masked_categorical(lg, wd, S, "sample", seed=seed, counter=9, env_id_base=base)
    a, lp = a.cpu().numpy(), lp.cpu().numpy()
    assert (a[~has] == 0).all() and (lp[~has] == 0).all()
    assert legal[np.arange(B)[has], a[has]].all()
    assert np.abs(lp[has] - ref[has, a[has]]).max() < 1e-5
    u = (philox_ref.draw(seed, base + np.arange(B, dtype=np.int64), 3, np.full(B, 9)).astype(np.int64) >> 8) / 16777216.0
    mx = np.where(legal, logits, -np.inf).max(1, keepdims=True)
    p = np.where(legal, np.exp(np.where(legal, logits.astype(np.float64) - np.where(np.isfinite(mx), mx, 0.0), 0.0)), 0.0)
    cdf = np.cumsum(p, axis=1)
    tot = cdf[:, -1]
    tgt = u * tot
    hi = cdf[np.arange(B), a]
    lo = hi - p[np.arange(B), a]
    slack = 1e-5 * np.maximum(tot, 1e-30)
    inside = (tgt >= lo - slack) & (tgt < hi + slack)
    assert inside[has].all()


def test_masked_categorical_distribution_and_counters():
    from gymothelloenv_b200 import masked_categorical
    S, B = 8, 200000
    rng = np.random.RandomState(0)
    row = (rng.randn(64) * 2).astype(np.float32)
    leg = np.zeros(64, dtype=bool)
    leg[[3, 17, 18, 40, 63]] = True
    word = np.uint64(sum(1 << i for i in np.nonzero(leg)[0]))
    lg = torch.from_numpy(np.tile(row, (B, 1))).cuda()
    wd = torch.from_numpy(np.full((1, B), word, dtype=np.uint64).view(np.int64)).cuda()
    a1, _ = masked_categorical(lg, wd, S, "sample", seed=3, counter=0)
    a2, _ = masked_categorical(lg, wd, S, "sample", seed=3, counter=1)
    a1b, _ = masked_categorical(lg, wd, S, "sample", seed=3, counter=0)
    assert torch.equal(a1, a1b) and not torch.equal(a1, a2)
    p = np.exp(row[leg] - row[leg].max())
    p /= p.sum()
    freq = np.bincount(a1.cpu().numpy(), minlength=64)[leg] / B
    assert np.abs(freq - p).max() < 5 * np.sqrt(0.25 / B) + 1e-3
    assert np.bincount(a1.cpu().numpy(), minlength=64)[~leg].sum() == 0


def test_sampler_drives_a_self_play_round_on_device():
    """A torch 'network' + LegalMoveSampler as the policy of PPOEnvs: the actions the env
    records are legal choices and the stored log-probs are the sampler's."""
    from gymothelloenv_b200 import PPOEnvs, LegalMoveSampler
    S, B = 8, 4096
    torch.manual_seed(0)
    net = torch.nn.Sequential(torch.nn.Flatten(), torch.nn.Linear(4 * S * S, S * S)).cuda()

    class Pol(object):
        def __init__(self):
            self.sampler = None

        def act(self, obs, hxs, masks, choices):
            logits = net(obs)
            a, lp = self.sampler.sample(logits)
            return logits.mean(1), a.unsqueeze(1), lp, torch.zeros_like(lp)

    pol = Pol()
    envs = PPOEnvs(B, None, pol, "cuda:0", board_size=S, num_disk_as_reward=True, seed=11)
    pol.sampler = LegalMoveSampler(envs.vec)
    envs.reset()
    assert set(np.unique(envs.protagonist.cpu().numpy())) == {-1, 1}
    n_tuples = 0
    for _ in range(40):
        obs, actions, rewards, dones, infos, outputs, masks, bad = envs.step(None)
        learn = (infos.type == 0) & (infos.choices_mask.any(1))
        idx = actions.long().squeeze(1)
        assert bool(infos.choices_mask[torch.arange(B, device="cuda"), idx][learn].all())
        assert bool((outputs[learn, 1] <= 0).all())
        n_tuples += int(learn.sum())
        if bool(dones.all()):
            break
    assert bool(dones.all()) and n_tuples > 25 * B
    r = rewards.squeeze(1)
    assert bool((r[infos.type == 1] == 0).all())
