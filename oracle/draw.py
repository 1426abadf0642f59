"""CPU restatement of the on-device batch draw (``mneslam_b200/csrc/assemble.cu``: ``draw_rays_kernel``).  TEST INFRASTRUCTURE ONLY.

What the reference does (``model/keyframe.py:98``, ``mp_slam/mapper.py:137-139``): ``random.sample(range(n), k)`` -- k DISTINCT
rows, uniformly.  Python's Mersenne-Twister stream cannot be reproduced on the device, so there is nothing to pin bit for bit
against the reference; what must hold is the PROPERTY (distinct rows inside the range, a fresh set per key), and the device
kernel must compute exactly the function stated here: element i of a draw is ``pi(i)`` for a keyed bijection ``pi`` of
``[0, n)`` -- a 4-round balanced Feistel network over the next even number of bits >= log2(n), walked along its cycle until
the value lands inside ``[0, n)``.  Pure Python integers; small cases only.
"""

M32 = 0xFFFFFFFF
TAG_KEYFRAMES = 0
TAG_CURRENT = 0x51ed270b


def mix32(x):
    x &= M32
    x ^= x >> 16
    x = (x * 0x7feb352d) & M32
    x ^= x >> 15
    x = (x * 0x846ca68b) & M32
    x ^= x >> 16
    return x


def draw_key(key, tag, n):
    """Round keys and the half width of the Feistel network for a population of ``n`` rows."""
    bits = 2
    while bits < 62 and (1 << bits) < n:
        bits += 2
    lo, hi = key & M32, (key >> 32) & M32
    ks = [mix32(lo ^ mix32((hi + 0x9e3779b9 * (r + 1) + tag) & M32)) for r in range(4)]
    return ks, bits // 2


def permute_index(i, n, ks, half_bits):
    mask = (1 << half_bits) - 1
    x = i
    while True:
        left, right = x >> half_bits, x & mask
        for r in range(4):
            f = mix32((right + ks[r]) & M32) & mask
            left, right = right, left ^ f
        x = (left << half_bits) | right
        if x < n:
            return x


def draw_rows(key, n_stored, n_global, n_current=0, n_cur=0):
    """Rows of the batch: ``n_global`` of ``n_stored`` keyframe rays, then ``n_cur`` of ``n_current`` current-frame pixels."""
    ks, hb = draw_key(key, TAG_KEYFRAMES, n_stored)
    rows = [permute_index(i, n_stored, ks, hb) for i in range(n_global)]
    if n_cur:
        ks, hb = draw_key(key, TAG_CURRENT, n_current)
        rows += [permute_index(i, n_current, ks, hb) for i in range(n_cur)]
    return rows
