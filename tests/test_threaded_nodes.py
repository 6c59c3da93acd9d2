"""Host logic of cpu-raytracer_b200/csrc/threaded_nodes.h, checked on the CPU through a g++ harness.

build_threaded_nodes lays a BLAS out in the order a traversal visits it, with a skip link per record, so
that the device walks it without a stack (traverse.cuh: blas_shadow; blas_trace_threaded under
RT_OCTANT_THREADED). What has to hold, and is checked here by walking both forms with the same
deterministic "is this box hit" predicate:

* shadow order (octant -1): the stackless walk reaches the same set of triangles as the reference's stack
  traversal (BottomLevelBVH.cpp:398-437) — an any-hit answer depends on nothing else;
* octant orders 0..7: the stackless walk tests the same boxes and the same triangles IN THE SAME SEQUENCE as
  the reference's closest-hit stack traversal (BottomLevelBVH.cpp:355-396 with BVHNode.h:30-40) for a ray
  of that octant, and a long leaf's box is tested once.
"""
import ctypes as C
import os
import zlib
import subprocess

import numpy as np
import pytest

from conftest import REPO, Golden

NODE = np.dtype([("min", "<f4", 3), ("max", "<f4", 3), ("left_or_first", "<i4"), ("count", "<i4")])


@pytest.fixture(scope="module")
def harness(tmp_path_factory):
    so = tmp_path_factory.mktemp("threaded") / "harness.so"
    subprocess.run(["g++", "-std=c++17", "-O1", "-fPIC", "-shared", "-Wall", "-Werror",
                    "-I", os.path.join(REPO, "cpu-raytracer_b200", "csrc"),
                    os.path.join(REPO, "tests", "threaded_nodes_harness.cpp"), "-o", str(so)], check=True)
    lib = C.CDLL(str(so))
    lib.threaded_build.restype = C.c_int
    lib.threaded_build.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int]

    def build(nodes, octant):
        nodes = np.ascontiguousarray(nodes)
        triangles = int(sum(n["count"] & 0x3FFFFFFF for n in nodes))
        out = np.zeros(len(nodes) + triangles // 63 + 8, dtype=NODE)
        n = lib.threaded_build(nodes.ctypes.data, len(nodes), octant, out.ctypes.data, len(out))
        return n, out[:max(n, 0)]
    return build


def blas_trees():
    """(name, nodes) for every BLAS of two fixtures + a synthetic tree with leaves beyond 63 triangles."""
    trees = []
    for name in ("dynamic_b3_aniso_f0", "dielectrics_b8_ewa"):
        g = Golden(name)
        nodes = g.pack.a["s_nodes"].view(NODE)
        for b, (node_offset, node_count, _, _, _) in enumerate(g.pack.a["s_blas"].view("<i4").reshape(-1, 5)):
            trees.append(("%s/blas%d" % (name, b), nodes[node_offset:node_offset + node_count].copy()))
    trees.append(("synthetic_long_leaves", synthetic_tree()))
    return trees


def synthetic_tree(seed=7):
    """A random binary tree in the reference's array form (root 0, node 1 unused, children in pairs) whose leaves hold
    1..200 triangles, every split axis, flattened triangle ranges."""
    rng = np.random.default_rng(seed)
    nodes = np.zeros(2 + 2 * 40, dtype=NODE)
    next_free, first_triangle = 2, 0
    todo = [0]
    while todo:
        n = todo.pop()
        nodes[n]["min"] = rng.uniform(-1, 0, 3)
        nodes[n]["max"] = rng.uniform(0, 1, 3)
        if next_free + 2 <= len(nodes) and (n == 0 or rng.random() < 0.7):
            nodes[n]["left_or_first"] = next_free
            nodes[n]["count"] = np.array(int(rng.integers(1, 4)) << 30, dtype=np.uint32).view(np.int32)   # split axis + 1
            todo += [next_free, next_free + 1]
            next_free += 2
        else:
            count = int(rng.choice([1, 2, 5, 63, 64, 126, 127, 200]))
            nodes[n]["left_or_first"] = first_triangle
            nodes[n]["count"] = count
            first_triangle += count
    return nodes[:next_free]


def box_key(n):
    return n["min"].tobytes() + n["max"].tobytes()


def predicate(seed):
    """Deterministic stand-in for AABB::intersect: a function of the box bytes only, so both walks agree on it."""
    if seed is None:
        return lambda key: True
    return lambda key: (zlib.crc32(key, seed) & 3) != 0        # not hash(): that one differs from process to process


def stack_walk(nodes, hit, left_first):
    """The reference's traversal: pop, test the box, leaf -> triangles, inner -> push far then near."""
    events, stack = [], [0]
    while stack:
        n = nodes[stack.pop()]
        key = box_key(n)
        events.append(("box", key))
        if not hit(key):
            continue
        count = int(n["count"]) & 0x3FFFFFFF
        left = int(n["left_or_first"])
        if count > 0:
            events += [("tri", i) for i in range(left, left + count)]
        elif left_first(int(n["count"])):
            stack += [left + 1, left]
        else:
            stack += [left, left + 1]
    return events


def threaded_walk(records, hit, octant_copy=False):
    """traverse.cuh: blas_shadow / (octant_copy) blas_trace_threaded, in Python."""
    events, index, steps = [], 0, 0
    while index >= 0:
        r = records[index]
        nxt, leaf = int(r["left_or_first"]), int(r["count"])
        ok = leaf < 0
        if not ok:
            key = box_key(r)
            events.append(("box", key))
            ok = hit(key)
        if ok:
            if leaf != 0:
                first, count = (leaf & (0x3FFFFFFF if octant_copy else 0x7FFFFFFF)) >> 6, leaf & 63
                events += [("tri", i) for i in range(first, first + count)]
                if octant_copy and leaf & 0x40000000:
                    nxt = index + 1
            else:
                nxt = index + 1
        index = nxt
        steps += 1
        assert steps <= len(records), "skip links loop"
    return events


def octant_left_first(octant):
    """BVHNode::should_visit_left_first (BVHNode.h:30-40) for a direction of that octant; axis code 0 falls to z."""
    positive = {1: bool(octant & 1), 2: bool(octant & 2), 3: bool(octant & 4), 0: bool(octant & 4)}
    return lambda count: positive[(count >> 30) & 3]


TREES = None


def trees():
    global TREES
    if TREES is None:
        TREES = blas_trees()
    return TREES


def test_octant_copies_replay_the_stack_traversal(harness):
    for name, nodes in trees():
        for octant in range(8):
            n, records = harness(nodes, octant)
            assert n > 0, name
            for seed in (None, 1, 2, 3, 4, 5, 6):
                want = stack_walk(nodes, predicate(seed), octant_left_first(octant))
                got = threaded_walk(records, predicate(seed), octant_copy=True)
                assert got == want, (name, octant, seed)


def test_naive_order_is_octant_seven(harness):
    """BVH_TRAVERSE_TREE_NAIVE: the left child first at every node."""
    for name, nodes in trees():
        _, records = harness(nodes, 7)
        assert threaded_walk(records, predicate(3), octant_copy=True) == stack_walk(nodes, predicate(3), lambda count: True), name


def test_shadow_copy_reaches_the_same_triangles(harness):
    for name, nodes in trees():
        n, records = harness(nodes, -1)
        assert n > 0, name
        assert not (records["count"] < 0).any()          # no "skip the box test" records in the shadow copy
        for seed in (None, 1, 2):
            want = stack_walk(nodes, predicate(seed), octant_left_first(0))
            got = threaded_walk(records, predicate(seed))
            assert sorted(e for e in got if e[0] == "tri") == sorted(e for e in want if e[0] == "tri"), (name, seed)
            # right child first: the same boxes, the other way round
            assert [e for e in got if e[0] == "box"][:1] == [e for e in want if e[0] == "box"][:1]


def test_every_copy_has_the_same_number_of_records(harness):
    """b200rt.cu places copy k of a BLAS at shadow_node_offset + (1 + k) * records_per_copy."""
    for name, nodes in trees():
        sizes = {harness(nodes, octant)[0] for octant in range(-1, 8)}
        assert len(sizes) == 1, (name, sizes)


def test_long_leaves_are_chained(harness):
    nodes = synthetic_tree()
    _, records = harness(nodes, 0)
    continued = records["count"] < 0
    assert continued.any()
    leaf = records["count"] != 0
    assert ((records["count"][leaf] & 63) >= 1).all()
    # a continuation record carries its leaf's box and follows the record it continues, which says "goes on" (bit 30);
    # every record of a chain links past the whole chain: a missed box skips all of it
    for i in np.flatnonzero(continued):
        assert records[i - 1]["count"] != 0 and records[i - 1]["count"] & 0x40000000
        assert records[i - 1]["left_or_first"] == records[i]["left_or_first"] and (records[i]["left_or_first"] == -1 or records[i]["left_or_first"] > i)
        assert box_key(records[i]) == box_key(records[i - 1])
    _, shadow = harness(nodes, -1)                     # the any-hit copy: plain links from record to record, no flags
    chained = [i for i in range(len(shadow) - 1) if shadow[i]["count"] != 0 and box_key(shadow[i]) == box_key(shadow[i + 1])]
    assert chained and all(shadow[i]["left_or_first"] == i + 1 for i in chained)


def test_broken_trees_are_refused(harness):
    nodes = synthetic_tree()
    bad = nodes.copy()
    bad[0]["left_or_first"] = len(nodes)                 # child out of bounds
    assert harness(bad, 0)[0] == -1
    loop = nodes.copy()
    loop[int(nodes[0]["left_or_first"])]["left_or_first"] = int(nodes[0]["left_or_first"])   # a child that is its own parent
    loop[int(nodes[0]["left_or_first"])]["count"] = 1 << 30
    assert harness(loop, -1)[0] == -1
