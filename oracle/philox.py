"""TEST INFRASTRUCTURE (oracle) -- not part of the product path.

NumPy restatement of the counter-based Philox4x32-10 generator used for the
prenet dropout masks in "perf mode" (SURVEY.md §7 'Injected randomness').  The
reference draws these bits from TF's stateful RNG (`tf.layers.dropout(...,
training=True)`, TacotronModel/modules/Decoder.py:350-351) which cannot be
reproduced; both sides of the parity test therefore use this keyed generator:

    counter = (step, batch_row, layer, unit >> 7), key = (seed_lo, seed_hi)
    keep(unit) = bit (unit & 127) of the 128-bit output block

Known-answer vectors from the Random123 distribution are checked in
tests/test_oracle_philox.py.
"""
import numpy as np

_M0 = np.uint64(0xD2511F53)
_M1 = np.uint64(0xCD9E8D57)
_W0 = 0x9E3779B9
_W1 = 0xBB67AE85
_MASK = np.uint64(0xFFFFFFFF)


def philox4x32_10(c0, c1, c2, c3, k0, k1):
    """Vectorised Philox4x32-10; all args broadcastable uint32 arrays. Returns 4 uint32 arrays."""
    c0 = np.asarray(c0, np.uint64) & _MASK
    c1 = np.asarray(c1, np.uint64) & _MASK
    c2 = np.asarray(c2, np.uint64) & _MASK
    c3 = np.asarray(c3, np.uint64) & _MASK
    c0, c1, c2, c3 = np.broadcast_arrays(c0, c1, c2, c3)
    k0 = int(k0) & 0xFFFFFFFF
    k1 = int(k1) & 0xFFFFFFFF
    for _ in range(10):
        p0 = _M0 * c0
        p1 = _M1 * c2
        hi0, lo0 = p0 >> np.uint64(32), p0 & _MASK
        hi1, lo1 = p1 >> np.uint64(32), p1 & _MASK
        c0, c1, c2, c3 = (hi1 ^ c1 ^ np.uint64(k0)), lo1, (hi0 ^ c3 ^ np.uint64(k1)), lo0
        k0 = (k0 + _W0) & 0xFFFFFFFF
        k1 = (k1 + _W1) & 0xFFFFFFFF
    return (c0.astype(np.uint32), c1.astype(np.uint32), c2.astype(np.uint32), c3.astype(np.uint32))


def prenet_keep_mask(seed, step, batch, units, layer):
    """uint8 [batch, units] keep bits for one decoder step and prenet layer."""
    b = np.arange(batch, dtype=np.uint64)[:, None]
    u = np.arange(units, dtype=np.uint64)[None, :]
    word = u >> np.uint64(7)
    o = philox4x32_10(np.uint64(step), b, np.uint64(layer), word, seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF)
    bit = (u & np.uint64(127)).astype(np.int64)
    lane = bit >> 5
    sh = (bit & 31).astype(np.uint32)
    sel = np.where(lane == 0, o[0], np.where(lane == 1, o[1], np.where(lane == 2, o[2], o[3])))
    return ((sel >> sh) & np.uint32(1)).astype(np.uint8)


def prenet_masks(seed, steps, batch, units=256, layers=2, step0=0):
    """uint8 [steps, batch, layers, units]."""
    out = np.empty((steps, batch, layers, units), np.uint8)
    for s in range(steps):
        for l in range(layers):
            out[s, :, l, :] = prenet_keep_mask(seed, step0 + s, batch, units, l)
    return out


TRAIN_TAG = 0x54524E31  # "TRN1": 4th counter word of the training-mode mask streams


def keep_threshold(p):
    """uint32 threshold t with keep <=> word < t  (P[keep] = t / 2^32)."""
    return min(int(p * 4294967296.0), 0xFFFFFFFF)


def keep_mask(seed, stream, n, p, offset=0):
    """Training-mode Bernoulli(p) keep bits (encoder / postnet dropout, zoneout masks; the reference
    draws them from TF's stateful RNG, Encoder.py:135-136,167): element ``idx`` of mask stream
    ``stream`` uses word ``idx & 3`` of Philox4x32-10(counter=(idx >> 2, stream, idx >> 34, TRAIN_TAG),
    key=seed).  Returns uint8 [n] for idx in [offset, offset + n)."""
    idx = np.arange(offset, offset + n, dtype=np.uint64)
    blk = idx >> np.uint64(2)
    o = philox4x32_10(blk & _MASK, np.uint64(stream), blk >> np.uint64(32), np.uint64(TRAIN_TAG),
                      seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF)
    lane = (idx & np.uint64(3)).astype(np.int64)
    word = np.where(lane == 0, o[0], np.where(lane == 1, o[1], np.where(lane == 2, o[2], o[3])))
    return (word < np.uint32(keep_threshold(p))).astype(np.uint8)
