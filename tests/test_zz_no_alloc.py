"""pg_her_sample_static (include/panda_b200.h): the HER sample() whose per-call scratch lives in a file-scope static
__device__ array of the library, for callers that must not allocate.  Called through the C ABI and compared with the existing
entry point pg_her_sample on the same MT19937 state: the batches are bit-identical (index, gather and reward kernels are the same
code; only where their scratch lives differs), on two calls in a row with new transitions stored in between; a second buffer
whose first sample is the static call covers the valid-index compaction over the static block counts, an empty buffer the status
word, and an oversize batch the refusal.  Collected last on purpose."""
import ctypes

import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

MAX_BATCH = 65536  # PG_HER_STATIC_MAX_BATCH


def _mt_state(seed):
    st = np.random.RandomState(seed).get_state()
    return np.ascontiguousarray(st[1], dtype=np.uint32), int(st[2])


def _set_state(lib, check, h, seed):
    key, pos = _mt_state(seed)
    check(lib.pg_her_set_mt_state(h, key.ctypes.data_as(ctypes.c_void_p), pos))


def _get_state(lib, check, h):
    key = np.zeros(624, np.uint32); pos = ctypes.c_int()
    check(lib.pg_her_get_mt_state(h, key.ctypes.data_as(ctypes.c_void_p), ctypes.byref(pos)))
    return key, pos.value


def test_static_scratch_sample_matches_pg_her_sample():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from panda_gym_b200 import _lib
    from panda_gym_b200.vec_env import PandaVecEnv
    from panda_gym_b200.her import HerReplayBuffer
    lib, check = _lib.load(), _lib.check
    E = 16
    env = PandaVecEnv("PandaReachAO-v3", n_envs=E, seed=11, scenario="reachao3", max_episode_steps=12)
    her = HerReplayBuffer(E * 64, env, n_sampled_goal=4, max_batch_size=4096, rng="numpy")
    dev = env.batch.device
    stream = ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    rng = np.random.default_rng(5)
    obs = env.reset()

    def roll(steps):
        nonlocal obs
        for _ in range(steps):
            a = rng.uniform(-1, 1, (E, 7)).astype(np.float32)
            nobs, rew, dones, infos = env.step(a)
            her.add(obs, nobs, a, rew, dones, timeouts=np.zeros(E, bool))
            obs = nobs

    roll(30)
    rs = (2 * 56 + 7 + 16 + 15) // 16 * 16
    for call, (B, seed) in enumerate(((1000, 3), (4096, 9))):  # two calls in a row: the array keeps the first call's contents
        if call == 1:
            roll(3)  # new transitions between the two rounds (pg_her_sample below rebuilds the valid list)
        # existing entry point
        out = _lib.HerBatch()
        _set_state(lib, check, her.h, seed)
        check(lib.pg_her_sample(her.h, env.batch.h, B, 1, 0, 0, ctypes.byref(out), stream))
        torch.cuda.synchronize(dev)
        assert int(out.row_stride) == rs
        ref = her._view(out.observation, (B, rs), rs).clone()
        ref_state = _get_state(lib, check, her.h)
        # new entry point: caller-owned output block, poisoned first
        got = torch.full((B, rs), float("nan"), device=dev)
        status = torch.full((1,), 7, dtype=torch.int32, device=dev)
        _set_state(lib, check, her.h, seed)
        torch.cuda.synchronize(dev)
        check(lib.pg_her_sample_static(her.h, env.batch.h, B, 1, 0, 0, ctypes.c_void_p(got.data_ptr()),
                                       ctypes.c_void_p(status.data_ptr()), stream))
        torch.cuda.synchronize(dev)
        assert int(status.item()) == 0
        used = 2 * 56 + 12 + 7 + 2  # observation .. done: the columns pg_her_batch exposes
        assert torch.equal(got[:, :used], ref[:, :used]), "static-scratch batch differs from pg_her_sample's (call %d)" % call
        got_state = _get_state(lib, check, her.h)
        assert np.array_equal(got_state[0], ref_state[0]) and got_state[1] == ref_state[1], "MT19937 state must advance alike"

    # the compaction over the static block counts: a fresh buffer whose FIRST sample is the static call
    her2 = HerReplayBuffer(E * 64, env, n_sampled_goal=4, max_batch_size=4096, rng="numpy")
    her1 = her
    obs = env.reset()
    for _ in range(30):
        a = rng.uniform(-1, 1, (E, 7)).astype(np.float32)
        nobs, rew, dones, infos = env.step(a)
        her2.add(obs, nobs, a, rew, dones, timeouts=np.zeros(E, bool))
        obs = nobs
    B = 2048
    got = torch.full((B, rs), float("nan"), device=dev)
    status = torch.full((1,), 7, dtype=torch.int32, device=dev)
    _set_state(lib, check, her2.h, 21)
    check(lib.pg_her_sample_static(her2.h, env.batch.h, B, 1, 0, 0, ctypes.c_void_p(got.data_ptr()), ctypes.c_void_p(status.data_ptr()), stream))
    torch.cuda.synchronize(dev)
    out = _lib.HerBatch()
    _set_state(lib, check, her2.h, 21)
    check(lib.pg_her_sample(her2.h, env.batch.h, B, 1, 0, 0, ctypes.byref(out), stream))
    torch.cuda.synchronize(dev)
    ref = her2._view(out.observation, (B, rs), rs)
    used = 2 * 56 + 12 + 7 + 2
    assert int(status.item()) == 0 and torch.equal(got[:, :used], ref[:, :used])

    # an empty buffer reports status 1 and leaves the caller's block alone; the next call clears the status again
    her3 = HerReplayBuffer(E * 64, env, n_sampled_goal=4, max_batch_size=512, rng="numpy")
    blk = torch.full((64, rs), 5.0, device=dev)
    _set_state(lib, check, her3.h, 1)
    check(lib.pg_her_sample_static(her3.h, env.batch.h, 64, 1, 0, 0, ctypes.c_void_p(blk.data_ptr()), ctypes.c_void_p(status.data_ptr()), stream))
    torch.cuda.synchronize(dev)
    assert int(status.item()) == 1 and bool((blk == 5.0).all())
    _set_state(lib, check, her1.h, 2)
    check(lib.pg_her_sample_static(her1.h, env.batch.h, 64, 1, 0, 0, ctypes.c_void_p(blk.data_ptr()), ctypes.c_void_p(status.data_ptr()), stream))
    torch.cuda.synchronize(dev)
    assert int(status.item()) == 0 and bool(torch.isfinite(blk[:, :used]).all())

    # sizes beyond the array are refused with a clear message (nothing is launched)
    rc = lib.pg_her_sample_static(her1.h, env.batch.h, MAX_BATCH + 1, 1, 0, 0, ctypes.c_void_p(blk.data_ptr()), None, stream)
    assert rc != 0 and b"PG_HER_STATIC_MAX_BATCH" in lib.pg_last_error()
    for b in (her1, her2, her3):
        b.close()
    env.close()
