"""Host logic of the four-wide any-hit tree (cpu-raytracer_b200/csrc/threaded_nodes.h: build_wide_nodes) and of the tree
validator (check_tree), on the CPU through the g++ harness.

The wide walk (traverse.cuh: blas_shadow_wide, -DRT_WIDE_SHADOW) skips the box test of every inner node whose children
were pulled up into its parent's record. That cannot change "which leaves does the ray reach" as long as (a) a skipped
node's box contains the boxes tested below it — checked here on the stored floats of every fixture tree and of the two
stand-in halls — and (b) the slab test is monotone in the box, which it is for regular rays (finite origin and 1/d):
checked here by replaying AABB::intersect (AABB.cpp:38-52, float32) over both trees for random rays and comparing the
sets of leaves reached.
"""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

from conftest import REPO
from test_threaded_nodes import NODE, blas_trees

F = np.float32


@pytest.fixture(scope="module")
def lib(tmp_path_factory):
    so = tmp_path_factory.mktemp("wide") / "harness.so"
    subprocess.run(["g++", "-std=c++17", "-O1", "-fPIC", "-shared", "-Wall", "-Werror", "-I", os.path.join(REPO, "cpu-raytracer_b200", "csrc"),
                    os.path.join(REPO, "tests", "threaded_nodes_harness.cpp"), "-o", str(so)], check=True)
    lib = C.CDLL(str(so))
    lib.wide_build.restype = C.c_int
    lib.wide_build.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int)]
    lib.tree_check.restype = C.c_int
    lib.tree_check.argtypes = [C.c_void_p, C.c_int, C.c_int, C.POINTER(C.c_int), C.c_char_p, C.c_int]
    return lib


def build_wide(lib, nodes):
    nodes = np.ascontiguousarray(nodes)
    out = np.zeros((len(nodes) + 1, 32), dtype=F)
    stack, escaped = C.c_int(), C.c_int()
    n = lib.wide_build(nodes.ctypes.data, len(nodes), out.ctypes.data, len(out), C.byref(stack), C.byref(escaped))
    return n, out[:max(n, 0)], stack.value, escaped.value


def slab_hit(lo, hi, o, inv_d, max_distance):
    """AABB::intersect in float32 with the reference's operations (regular ray: no NaN can arise)."""
    t0, t1 = (lo - o) * inv_d, (hi - o) * inv_d
    t_near = max(F(0.005), min(t0[0], t1[0]), min(t0[1], t1[1]), min(t0[2], t1[2]))
    t_far = min(max_distance, max(t0[0], t1[0]), max(t0[1], t1[1]), max(t0[2], t1[2]))
    return t_near < t_far


def leaves_binary(nodes, o, inv_d, max_distance):
    reached, stack = set(), [0]
    while stack:
        n = nodes[stack.pop()]
        if not slab_hit(n["min"], n["max"], o, inv_d, max_distance):
            continue
        count = int(n["count"]) & 0x3FFFFFFF
        if count:
            reached.add((int(n["left_or_first"]), count))
        else:
            stack += [int(n["left_or_first"]), int(n["left_or_first"]) + 1]
    return reached


def leaves_wide(records, o, inv_d, max_distance):
    reached, stack, deepest = set(), [0], 1
    while stack:
        r = records[stack.pop()]
        links, counts = r[24:28].view(np.int32), r[28:32].view(np.int32)
        for k in range(4):
            if counts[k] < 0:
                continue
            lo, hi = np.array([r[k], r[4 + k], r[8 + k]]), np.array([r[12 + k], r[16 + k], r[20 + k]])
            if not slab_hit(lo, hi, o, inv_d, max_distance):
                continue
            if counts[k] == 0:
                stack.append(int(links[k]))
            else:
                reached.add((int(links[k]), int(counts[k])))
        deepest = max(deepest, len(stack))
    return reached, deepest


@pytest.mark.parametrize("name,nodes", blas_trees(), ids=[t[0] for t in blas_trees()])
def test_wide_walk_reaches_the_binary_walks_leaves(lib, name, nodes):
    n, records, stack_entries, escaped = build_wide(lib, nodes)
    if name.startswith("synthetic"):                    # random boxes: the root's box does not contain its children, the builder must refuse
        assert n == -1
        return
    assert n > 0 and escaped == 0                       # the builder's boxes always contain their children's
    assert n <= (len(nodes) + 1) // 2 + 1               # about half the records of the binary tree
    rng = np.random.default_rng(11)
    lo, hi = nodes[0]["min"], nodes[0]["max"]
    worst = 0
    for _ in range(60):
        o = (lo + (hi - lo) * rng.uniform(-0.3, 1.3, 3)).astype(F)
        d = rng.normal(size=3).astype(F)
        d /= np.linalg.norm(d)
        with np.errstate(divide="ignore"):
            inv_d = (F(1.0) / d).astype(F)
        max_distance = F(rng.choice([np.inf, 0.5 * float(np.linalg.norm(hi - lo))]))
        wide, deepest = leaves_wide(records, o, inv_d, max_distance)
        assert wide == leaves_binary(nodes, o, inv_d, max_distance)
        worst = max(worst, deepest)
    assert worst <= stack_entries                        # the stack bound the upload checks against RT_WIDE_STACK


def test_stand_in_halls_are_contained_and_fit_the_stack(lib):
    """The two big trees the benchmark runs on: no box escapes its parent's, and the pending-record bound fits RT_WIDE_STACK (48)."""
    from b200rt import standin
    from b200rt.host import HostScene
    for scene_file in (standin.sponza_standin(width=64, height=36), standin.sibenik_standin(width=64, height=36)):
        scene = HostScene(scene_file)
        scene.update(0.0, 0)
        pack = scene.pack()
        nodes = pack.a["s_nodes"].view(NODE)
        for node_offset, node_count, _, _, _ in pack.blas():
            n, _, stack_entries, escaped = build_wide(lib, nodes[node_offset:node_offset + node_count].copy())
            assert n > 0 and escaped == 0 and stack_entries <= 48


def test_check_tree_rejects_what_the_kernels_would_trust(lib):
    """a leaf whose count carries axis bits, an inner node with axis code 0 and bad children, a cycle."""
    def check(nodes, leaf_limit):
        nodes = np.ascontiguousarray(nodes)
        depth, why = C.c_int(), C.create_string_buffer(128)
        rc = lib.tree_check(nodes.ctypes.data, len(nodes), leaf_limit, C.byref(depth), why, 128)
        return rc, depth.value, why.value.decode()

    good = np.zeros(4, dtype=NODE)
    good[0]["left_or_first"], good[0]["count"] = 2, 1 << 30            # inner, split axis x, children 2 and 3
    good[2]["left_or_first"], good[2]["count"] = 0, 2
    good[3]["left_or_first"], good[3]["count"] = 2, 3
    good[1]["left_or_first"], good[1]["count"] = 12345, 77              # node 1 is never reached: anything goes
    assert check(good, 5)[:2] == (0, 1)

    bad = good.copy(); bad[3]["count"] = 0x40000003                    # leaf count with axis bits: ~1e9 loop iterations on the device
    assert check(bad, 5)[0] == -1 and "axis bits" in check(bad, 5)[2]
    bad = good.copy(); bad[3]["left_or_first"] = 3                     # leaf range past the triangles
    assert check(bad, 5)[0] == -1
    bad = good.copy(); bad[0]["count"] = 0; bad[0]["left_or_first"] = 3   # axis code 0 and a right child beyond the array
    assert check(bad, 5)[0] == -1 and "child" in check(bad, 5)[2]
    bad = good.copy(); bad[2]["count"] = 1 << 30; bad[2]["left_or_first"] = 2   # node 2 is its own child
    assert check(bad, 5)[0] == -1 and "tree" in check(bad, 5)[2]
