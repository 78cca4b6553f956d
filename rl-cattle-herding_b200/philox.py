"""Host mirror of the device RNG (csrc/chs_math.cuh: Philox4x32-10, Salmon et al. SC'11).

Resets draw from Philox keyed by the 64-bit seed with counter (global env id, episode index, draw index,
stream tag); this module reproduces those draws in numpy so that tests (and users who checkpoint) can predict
what a device reset will produce.  The reference itself draws from the process-global Mersenne Twister
(`random.randint`, `np.random.rand`: sb3_envs/BaseAviary.py:307,617,631), which is not reproducible per env.
"""
import numpy as np

M0, M1 = np.uint64(0xD2511F53), np.uint64(0xCD9E8D57)
W0, W1 = 0x9E3779B9, 0xBB67AE85
TAG_COW, TAG_NDRONES = 0, 1
MASK32 = np.uint64(0xFFFFFFFF)


def philox4x32_10(c0, c1, c2, c3, seed):
    """Vectorised over numpy arrays of uint32 counters.  Returns four uint32 arrays."""
    c = [np.asarray(x, dtype=np.uint64) & MASK32 for x in np.broadcast_arrays(c0, c1, c2, c3)]
    k0, k1 = int(seed) & 0xFFFFFFFF, (int(seed) >> 32) & 0xFFFFFFFF
    for _ in range(10):
        p0 = M0 * c[0]
        p1 = M1 * c[2]
        n0 = ((p1 >> np.uint64(32)) ^ c[1] ^ np.uint64(k0)) & MASK32
        n1 = p1 & MASK32
        n2 = ((p0 >> np.uint64(32)) ^ c[3] ^ np.uint64(k1)) & MASK32
        n3 = p0 & MASK32
        c = [n0, n1, n2, n3]
        k0, k1 = (k0 + W0) & 0xFFFFFFFF, (k1 + W1) & 0xFFFFFFFF
    return [x.astype(np.uint32) for x in c]


def u01_24(x):
    """24-bit uniform in [0,1), exactly representable in fp32 and fp64."""
    return (np.asarray(x, dtype=np.uint32) >> np.uint32(8)).astype(np.float64) / 16777216.0


def cow_uniforms(seed, global_env_id, episode, n_cattle):
    """(n_cattle, 2) uniforms: column 0 stands for the yaw draw (unused on the DYN path), column 1 for the
    initial-velocity angle draw of BaseAviary.py:617,631."""
    j = np.arange(n_cattle, dtype=np.uint64)
    r = philox4x32_10(np.uint64(global_env_id), np.uint64(episode), j, np.uint64(TAG_COW), seed)
    return np.stack([u01_24(r[0]), u01_24(r[1])], axis=1)


def num_drones(seed, global_env_id, episode, min_n, max_n, max_slots=12):
    """`random.randint(MIN, MAX)` of BaseAviary.py:307 as the device draws it."""
    r = philox4x32_10(np.uint64(global_env_id), np.uint64(episode), np.uint64(0), np.uint64(TAG_NDRONES), seed)
    n = int(min_n + int(float(u01_24(r[0])) * (max_n - min_n + 1)))
    return min(n, max_slots)
