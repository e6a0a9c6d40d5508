"""TEST INFRASTRUCTURE -- a deterministic stand-in for the PPO policy network.

``act`` has the call signature of the reference's ``Policy.act`` as used by
``PPOEnvs.step`` (envs.py:203; model.py:60): (obs, rnn_hxs, masks, choices) ->
(value, action, log_prob, hidden).  Every output is a pure function of (env index,
observation planes, legal choices), so the live reference (tests/golden/
make_golden_selfplay.py) and the device implementation can be driven identically and their
outputs compared bit for bit.  All returned floats are exactly representable in float32.
"""
import numpy as np

_MOD = 1000003


def obs_hash(obs, board_size):
    """obs: array-like (B, 4, S, S) of 0/1 -> int64[B] in [0, _MOD)"""
    o = np.asarray(obs).astype(np.int64).reshape(len(obs), -1)
    n = o.shape[1]
    w = (np.arange(n, dtype=np.int64) * 2654435761 + 12345) % _MOD
    return ((o * w[None, :]).sum(axis=1) + 7919 * np.arange(len(o), dtype=np.int64)) % _MOD


def decide(h, choices_i):
    """-> (action, v, logp, hidden) for hash h and an ascending list of legal actions"""
    h = int(h)
    a = int(choices_i[h % len(choices_i)]) if len(choices_i) else 0
    v = np.float32((h % 1024) / 64.0 - 8.0)
    lp = np.float32(-((h // 7) % 512) / 32.0)
    hid = np.float32(((h // 11) % 256) / 16.0)
    return a, v, lp, hid


class FakePolicy(object):
    def __init__(self, board_size, salt=0):
        self.board_size = int(board_size)
        self.salt = int(salt)             # teacher / student stand-ins differ by salt (0 = the self-play golden)
        self.calls = 0

    def act(self, obs, rnn_hxs, masks, choices, deterministic=False):
        if hasattr(obs, "detach"):
            obs = obs.detach().cpu().numpy()
        if not isinstance(choices, (list, tuple)):        # dense bool mask (B, N) from the device env
            m = choices.detach().cpu().numpy() if hasattr(choices, "detach") else np.asarray(choices)
            choices = [np.nonzero(row)[0].tolist() for row in m]
        self.calls += 1
        hs = (obs_hash(obs, self.board_size) + self.salt) % _MOD
        B = len(hs)
        act = np.zeros((B, 1), dtype=np.int64)
        v = np.zeros((B,), dtype=np.float32)
        lp = np.zeros((B,), dtype=np.float32)
        hid = np.zeros((B,), dtype=np.float32)
        for i in range(B):
            act[i, 0], v[i], lp[i], hid[i] = decide(hs[i], choices[i])
        return v, act, lp, hid
