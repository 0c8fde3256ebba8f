"""CPU models of the round-2 changes to the hand-written onesweep sort (csrc/build_sort.cu), checkable without a GPU:

* rt_match_digit8: the lanes that share my 8-bit digit, from eight ballots (one per bit) instead of MATCH.ANY; in the
  Morton kernel's histogram the invalid lanes of a partial warp carry 256 + lane and are masked out afterwards;
* the hybrid histogram: plain per-lane counts unless more than 8 lanes share the first lane's digit, then one count per
  group - the same totals either way;
* the windowed decoupled look-back: a tile reads LB_WINDOW predecessors' states, each load at its own moment, and consumes
  them in order until it meets an inclusive prefix or an unpublished entry.  Tiles run as coroutines that yield at every
  memory operation and a random scheduler interleaves them: every tile must end with the exact exclusive prefix of the
  tile counts before it, for any interleaving.

tests/test_build_gpu.py holds the kernels themselves to the oracle (sorted codes and permutation bit for bit)."""
import numpy as np
import pytest

LB_AGG, LB_INC, LB_VAL = 1 << 30, 2 << 30, (1 << 30) - 1
LB_WINDOW = 8


def match_digit8(d):
    """d: 32 lane values; returns the 32 peer masks of rt_match_digit8 (only the low 8 bits take part)"""
    d = np.asarray(d, np.int64)
    peers = np.full(32, 0xFFFFFFFF, np.int64)
    for b in range(8):
        bit = (d >> b) & 1
        m = int(sum(1 << i for i in range(32) if bit[i]))
        peers &= np.where(bit == 1, m, ~m & 0xFFFFFFFF)
    return peers


@pytest.mark.parametrize("seed", range(6))
def test_ballot_match_equals_groups_of_equal_digits(seed):
    rng = np.random.RandomState(seed)
    for spread in (256, 16, 2, 1):
        d = rng.randint(0, spread, 32)
        peers = match_digit8(d)
        for i in range(32):
            assert peers[i] == sum(1 << j for j in range(32) if d[j] == d[i])
    # a partial warp of the Morton kernel: invalid lanes carry 256 + lane; an invalid lane's LOW 8 bits may equal a valid digit
    nvalid = rng.randint(1, 32)
    d = np.r_[rng.randint(0, 256, nvalid), 256 + np.arange(nvalid, 32)]
    d[0] = nvalid if nvalid < 32 else d[0]  # provoke the collision: lane `nvalid` carries 256 + nvalid
    act = (1 << nvalid) - 1
    peers = match_digit8(d) & act
    for i in range(nvalid):
        assert peers[i] == sum(1 << j for j in range(nvalid) if d[j] == d[i])


@pytest.mark.parametrize("seed", range(4))
def test_hybrid_histogram_counts_every_valid_lane_once(seed):
    rng = np.random.RandomState(seed)
    for spread, nvalid in ((256, 32), (3, 32), (1, 32), (256, 5), (1, 5), (2, 20)):
        d = rng.randint(0, spread, 32)
        valid = np.arange(32) < nvalid
        hist = np.zeros(256, np.int64)
        first = 0
        same = sum(1 for i in range(32) if valid[i] and d[i] == d[first])
        if same > 8:  # aggregated path: the lowest lane of each group adds the group's size
            dd = np.where(valid, d, 256 + np.arange(32))
            peers = match_digit8(dd) & ((1 << nvalid) - 1)
            for i in range(32):
                if valid[i] and (int(peers[i]) & -int(peers[i])) == (1 << i):
                    hist[d[i]] += bin(int(peers[i])).count("1")
        else:
            for i in range(32):
                if valid[i]:
                    hist[d[i]] += 1
        assert np.array_equal(hist, np.bincount(d[:nvalid], minlength=256))


def _tile(t, counts, lookback, result):
    """one digit-thread of tile t, as in onesweep_kernel: publish, then walk back through windows of LB_WINDOW loads"""
    cnt = int(counts[t])
    lookback[t] = (LB_INC if t == 0 else LB_AGG) | cnt
    yield
    excl = 0
    if t > 0:
        tt = t - 1; done = False
        while not done:
            v = []
            for k in range(LB_WINDOW):  # the loads of one window happen one after the other, other tiles run in between
                v.append(lookback[tt - k] if tt - k >= 0 else LB_INC)
                yield
            for k in range(LB_WINDOW):
                if done or (v[k] >> 30) == 0:
                    break
                excl += v[k] & LB_VAL; tt -= 1
                done = (v[k] & LB_INC) != 0
        lookback[t] = LB_INC | (excl + cnt)
        yield
    result[t] = excl


@pytest.mark.parametrize("tiles,resident", [(1, 1), (2, 2), (9, 9), (40, 40), (245, 245), (245, 37), (600, 64)])
def test_windowed_lookback_gives_exact_prefixes_for_any_interleaving(tiles, resident):
    for seed in range(3):
        rng = np.random.RandomState(seed + tiles)
        counts = rng.randint(0, 4097, tiles)
        lookback = [0] * tiles
        result = [None] * tiles
        # tiles are handed out by ticket: at most `resident` run at once, the next ticket starts when a slot frees up
        nxt = 0; live = []
        while nxt < tiles or live:
            while nxt < tiles and len(live) < resident:
                live.append(_tile(nxt, counts, lookback, result)); nxt += 1
            g = live[rng.randint(len(live))]
            try:
                next(g)
            except StopIteration:
                live.remove(g)
        want = np.concatenate([[0], np.cumsum(counts)[:-1]])
        assert np.array_equal(np.array(result, np.int64), want)
        assert all((lookback[t] >> 30) == 2 and (lookback[t] & LB_VAL) == want[t] + counts[t] for t in range(tiles))
