"""CPU model of the round-2 builder's climb (csrc/build_tree.cu, bottom_up_warp_kernel) against the oracle's tree.

The kernel's protocol, restated in Python lane by lane: the 32 consecutive leaves of a warp first meet inside the warp
(a left child [l, s] looks into lane s + 1 - k0, where its right sibling must sit if it is finished; a right child announces
its parent), the survivors continue with the global protocol (first arrival deposits and leaves, second one unions and goes
on), served here in a random order to stand for any interleaving on the GPU.  The tree that comes out - every internal
node's leaf range, and the split position as the node's index - must be the tree the reference's top-down Karras search
builds (oracle/oracle.cpp, restating src/lbvh_builder.cu:103-322), for any arrival order, at sizes around the warp border,
with duplicate keys, and when every key is equal (the shipped reference).

This covers the host-checkable part of the algorithm without a GPU; tests/test_build_gpu.py holds the kernel itself to
the oracle bit for bit."""
import sys

import numpy as np
import pytest

import helpers as H
import oracle_lib as O
from cuda_ray_tracer_b200 import synth


def _clz32(x):
    return 32 - int(x).bit_length()


class Delta:
    """Delta<uint32_t> of build_tree.cu: common-prefix length of neighbours, index tie-break for equal keys, -1 outside"""

    def __init__(self, keys):
        self.k = [int(x) for x in keys]; self.n = len(keys)

    def adj(self, i):
        if i < 0 or i >= self.n - 1:
            return -1
        a, b = self.k[i], self.k[i + 1]
        if a == b:
            return 32 + _clz32(i ^ (i + 1))
        return _clz32(a ^ b)


def climb(keys, rng):
    """returns {s: (left_ref, right_ref)} with ref = ('L', k) or ('N', s), and the root ref"""
    n = len(keys); D = Delta(keys)
    nodes = {}; root = [None]; survivors = []
    for k0 in range(0, n, 32):
        klast = min(k0 + 31, n - 1)
        lanes = []
        for lane in range(32):
            k = k0 + lane
            lanes.append(dict(active=k < n, l=k, r=k, ref=("L", k), need=True, s=-1, side=0))
        while True:
            for st in lanes:  # (s, side) of every live lane whose range has changed
                if st["active"] and st["need"]:
                    if st["l"] == 0 and st["r"] == n - 1:
                        root[0] = st["ref"]; st["active"] = False
                    else:
                        dl, dr = D.adj(st["l"] - 1), D.adj(st["r"])
                        st["s"], st["side"] = (st["r"], 0) if dr > dl else (st["l"] - 1, 1)
                        st["need"] = False
            taken = {}
            for lane, st in enumerate(lanes):  # the left child looks into the lane where its sibling must sit
                if st["active"] and st["side"] == 0 and st["s"] + 1 <= klast:
                    want = st["s"] + 1 - k0
                    p = lanes[want]
                    tag = p["s"] if (p["active"] and p["side"] == 1) else -2
                    if tag == st["s"]:
                        assert p["l"] == st["s"] + 1, "a subtree must sit in the lane of its leftmost leaf"
                        taken[lane] = want
            if not taken:
                break
            assert len(set(taken.values())) == len(taken)
            for lane, want in taken.items():
                st, p = lanes[lane], lanes[want]
                assert st["s"] not in nodes
                nodes[st["s"]] = (st["ref"], p["ref"])
                st["r"] = p["r"]; st["ref"] = ("N", st["s"]); st["need"] = True
                p["active"] = False
        survivors += [st for st in lanes if st["active"]]
    # global protocol, arrivals in a random order
    flags = {}
    work = list(survivors)
    while work:
        st = work.pop(rng.randint(len(work)))
        dl, dr = D.adj(st["l"] - 1), D.adj(st["r"])
        s, side, mine = (st["r"], 0, st["l"]) if dr > dl else (st["l"] - 1, 1, st["r"])
        if s not in flags:
            flags[s] = (mine, st["ref"], side)  # deposit the half and leave
            continue
        other, oref, oside = flags[s]
        assert oside != side and s not in nodes
        nodes[s] = (st["ref"], oref) if side == 0 else (oref, st["ref"])
        if side == 0:
            st["r"] = other
        else:
            st["l"] = other
        st["ref"] = ("N", s)
        if st["l"] == 0 and st["r"] == n - 1:
            root[0] = st["ref"]
        else:
            work.append(st)
    return nodes, root[0]


def ranges_of_model(nodes, root, n):
    out = {}

    def rec(ref):
        if ref[0] == "L":
            return ref[1], ref[1]
        l0, r0 = rec(nodes[ref[1]][0])
        l1, r1 = rec(nodes[ref[1]][1])
        assert r0 == ref[1] and l1 == ref[1] + 1, "node index must be the split position"
        out[ref[1]] = (l0, r1)
        return l0, r1

    sys.setrecursionlimit(max(10000, 4 * n))
    rec(root)
    return out


def ranges_of_oracle(b):
    n = b["n"]; left, right = b["left"], b["right"]
    out = set()
    stack = [(0, False)]; rng_of = {}
    while stack:
        node, done = stack.pop()
        if node >= n - 1:
            rng_of[node] = (node - (n - 1), node - (n - 1)); continue
        if not done:
            stack.append((node, True)); stack.append((int(left[node]), False)); stack.append((int(right[node]), False))
        else:
            l = rng_of[int(left[node])]; r = rng_of[int(right[node])]
            assert l[1] + 1 == r[0]
            rng_of[node] = (l[0], r[1]); out.add((l[0], r[1], l[1]))  # (first leaf, last leaf, split)
    return out


@pytest.mark.parametrize("n", [2, 3, 31, 32, 33, 63, 64, 65, 97, 255, 256, 257, 700])
@pytest.mark.parametrize("mode", [O.BOUNDS_COMPUTED, O.BOUNDS_SHIPPED])
def test_warp_climb_model_builds_the_oracle_tree(n, mode):
    cfg, a = synth.soup(n, seed=100 + n, width=16, height=16, aa=0)
    if mode == O.BOUNDS_COMPUTED and n >= 64:  # duplicate keys: the second half repeats vertices of the first
        for t in a["tri"][:3]:
            t[n // 2:] = t[: n - n // 2]
    a = synth.finish_triangles(a)
    o = H.oracle_from_arrays(cfg, a)
    b = o.build(mode, 30)
    want = ranges_of_oracle(b)
    assert len(want) == n - 1
    for seed in range(3):  # three arrival orders of the global part
        nodes, root = climb(b["codes"], np.random.RandomState(seed))
        assert len(nodes) == n - 1 and root is not None
        got = ranges_of_model(nodes, root, n)
        assert {(l, r, s) for s, (l, r) in got.items()} == want
