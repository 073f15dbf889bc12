"""LBVH build invariants on the device-built tree (SURVEY §4 property tests): sorted Morton keys,
slot permutation, every primitive in exactly one leaf, parent boxes contain child boxes, node count,
huge-primitive handling, duplicate keys, rebuild determinism."""
import numpy as np
import pytest

import scenes

pytestmark = pytest.mark.gpu

LEAF = 0x80000000


def download(gs, n):
    st = gs.scene.stats()
    nn = int(st.n_nodes)
    nodes = np.zeros((nn, 16), np.float32)
    geom = np.zeros((max(n, 1), 4), np.float32)
    prim = np.zeros(max(n, 1), np.uint32)
    keys = np.zeros(max(n, 1), np.uint32)
    gs.scene.download(nodes.ctypes.data, geom.ctypes.data, prim.ctypes.data, keys.ctypes.data)
    return st, nodes, geom[:n], prim[:n], keys[:n]


def slot_box(world, gid):
    ki = int(world["kind_index"][gid])
    if ki >> 31:
        v = world["tri_verts"][ki & 0x7FFFFFFF].reshape(3, 3)
        return v.min(0), v.max(0)
    s = world["spheres"][ki]
    return s[:3] - s[3], s[:3] + s[3]


def check_tree(st, nodes, prim, keys, world):
    n_bvh = int(st.n_bvh_prims)
    n = len(prim)
    assert sorted(prim.tolist()) == list(range(n)), "slot_prim must be a permutation of the global ids"
    assert np.all(np.diff(keys[:n_bvh].astype(np.int64)) >= 0), "Morton keys must be sorted"
    assert np.all(keys[:n_bvh] < (1 << 30)) and np.all(keys[n_bvh:] == 0xFFFFFFFF)
    if n_bvh < 2:
        return
    assert int(st.n_nodes) == n_bvh - 1
    child = nodes[:, 12:14].copy().view(np.int32)
    seen = np.zeros(n_bvh, np.int32)
    visited_nodes = 0
    stack = [(0, None)]
    while stack:
        ni, box = stack.pop()
        visited_nodes += 1
        nd = nodes[ni]
        for c in (0, 1):
            if c == 0:
                lo = np.array([nd[0], nd[2], nd[8]]); hi = np.array([nd[1], nd[3], nd[9]])
            else:
                lo = np.array([nd[4], nd[6], nd[10]]); hi = np.array([nd[5], nd[7], nd[11]])
            assert np.all(lo <= hi)
            if box is not None:
                assert np.all(lo >= box[0]) and np.all(hi <= box[1]), "child box escapes its parent"
            ch = int(child[ni, c])
            if ch < 0:
                enc = ch & 0xFFFFFFFF
                first, cnt = enc & 0x07FFFFFF, ((enc >> 27) & 7) + 1
                for s in range(first, first + cnt):
                    seen[s] += 1
                    plo, phi = slot_box(world, int(prim[s]))
                    assert np.all(plo >= lo) and np.all(phi <= hi), "leaf box does not contain its primitive"
            else:
                stack.append((ch, (lo, hi)))
    assert np.all(seen == 1), "every primitive must be in exactly one leaf"
    assert visited_nodes <= n_bvh - 1


@pytest.mark.parametrize("n,ntri,ground", [(2, 0, False), (5, 3, False), (64, 16, True), (1000, 200, True), (50000, 0, False)])
@pytest.mark.parametrize("refit_mode", [0, 1, 2])
def test_tree_invariants(pkg, synth, n, ntri, ground, refit_mode):
    d = scenes.random_scene(pkg, synth, n, n_tris=ntri, seed=n, with_walls=n >= 64, with_ground=ground)
    world = scenes.flatten_world(d)
    gs = scenes.GpuScene(pkg, d, options={"wide": 0, "refit_mode": refit_mode})
    st, nodes, geom, prim, keys = download(gs, len(world["kind_index"]))
    assert int(st.n_prims) == len(world["kind_index"])
    assert int(st.n_huge) == (1 if ground else 0)
    check_tree(st, nodes, prim, keys, world)
    # slot geometry is the gathered world geometry
    sph = (world["kind_index"][prim] >> 31) == 0
    assert np.array_equal(geom[sph], world["spheres"][world["kind_index"][prim[sph]]])
    # scene bounds cover every BVH primitive
    lo = np.array(list(st.scene_lo)); hi = np.array(list(st.scene_hi))
    for s in range(0, int(st.n_bvh_prims), max(1, int(st.n_bvh_prims) // 200)):
        plo, phi = slot_box(world, int(prim[s]))
        assert np.all(plo >= lo) and np.all(phi <= hi)
    gs.close()


def test_duplicate_morton_keys(pkg, synth):
    """Many primitives in one Morton cell (identical centroids): the index tie-break must still give
    a valid tree, and traversal must still equal brute force."""
    c = np.zeros((300, 3), np.float32)
    c[:100] = [0.1, 0.1, 0.1]
    c[100:200] = [0.1, 0.1, 0.1000001]
    c[200:] = np.random.default_rng(0).uniform(-0.4, 0.4, (100, 3))
    r = np.linspace(0.01, 0.03, 300).astype(np.float32)
    d = scenes.SceneDesc(scenes.CONFIG_ROUGHS, scenes.CONFIG_METALS)
    g = d.add_spheres(c, r)
    d.add_instance(g, None, 0, 0, 0)
    world = scenes.flatten_world(d)
    gs = scenes.GpuScene(pkg, d, options={"wide": 0})
    st, nodes, geom, prim, keys = download(gs, 300)
    assert len(np.unique(keys)) < 250
    check_tree(st, nodes, prim, keys, world)
    gs.close()


def test_rebuild_is_deterministic_and_tracks_motion(pkg, synth):
    import torch
    centers, radii = synth.uniform_spheres(20000, 0.004, 5)
    d = scenes.SceneDesc(scenes.CONFIG_ROUGHS, scenes.CONFIG_METALS)
    g = d.add_spheres(centers, radii)
    d.add_instance(g, None, 0, 0, 0)
    gs = scenes.GpuScene(pkg, d, options={"wide": 0})
    st1, n1, g1, p1, k1 = download(gs, 20000)
    gs.scene.build()
    st2, n2, g2, p2, k2 = download(gs, 20000)
    assert np.array_equal(p1, p2) and np.array_equal(k1, k2) and np.array_equal(n1.view(np.uint32), n2.view(np.uint32))
    # move the particles in place (caller-owned buffer) and rebuild: the tree follows
    moved = centers + np.float32([0.0, 0.0, -0.05])
    gs.keep[2].copy_(torch.from_numpy(moved).cuda())
    gs.scene.build()
    d.groups[0]["a"] = moved
    st3, n3, g3, p3, k3 = download(gs, 20000)
    check_tree(st3, n3, p3, k3, scenes.flatten_world(d))
    assert abs(st3.scene_lo[2] - (st1.scene_lo[2] - 0.05)) < 1e-4
    assert st3.build_ms > 0
    gs.close()


def check_wide_tree(gs, world):
    """8-wide compressed BVH: every primitive referenced exactly once, every child's decoded box
    contains its primitives' boxes / its children's decoded boxes, internal children contiguous."""
    n = len(world["kind_index"])
    st = gs.scene.stats()
    n_bvh = int(st.n_bvh_prims)
    nn = gs.scene.download_wide()
    assert nn == int(st.n_nodes)
    wn = np.zeros((max(nn, 1), 20), np.float32)
    geom = np.zeros((max(n, 1), 4), np.float32)
    prim = np.zeros(max(n, 1), np.uint32)
    gs.scene.download_wide(wn.ctypes.data, geom.ctypes.data, prim.ctypes.data)
    prim = prim[:n]
    assert sorted(prim.tolist()) == list(range(n)), "re-packed slot_prim must be a permutation of the global ids"
    if n_bvh == 0:
        return
    wu = wn.view(np.uint32)
    seen = np.zeros(n, np.int32)
    node_seen = np.zeros(nn, np.int32)
    stack = [(0, [])]
    while stack:
        w, anc = stack.pop()
        node_seen[w] += 1
        org = wn[w, 0:3].astype(np.float64)
        eim = int(wu[w, 3])
        scale = np.array([2.0 ** (((eim >> (8 * a)) & 0xFF) - 127) for a in range(3)])
        imask = eim >> 24
        child_base, prim_base = int(wu[w, 4]), int(wu[w, 5])
        leafmask = int(wu[w, 6]) & 0xFF
        assert int(wu[w, 6]) >> 8 == 0 and (imask & leafmask) == 0
        q = wu[w, 8:20].copy().view(np.uint8).reshape(6, 8)  # qlo_x qlo_y qlo_z qhi_x qhi_y qhi_z
        rank = off = 0
        for k in range(8):
            lo = org + q[0:3, k] * scale
            hi = org + q[3:6, k] * scale
            if not ((imask | leafmask) >> k) & 1:
                assert np.all(q[0:3, k] == 255) and np.all(q[3:6, k] == 0), "empty slots carry an inverted box"
                continue
            assert np.all(lo <= hi)
            if (imask >> k) & 1:
                stack.append((child_base + rank, anc + [(lo, hi)]))   # internal children contiguous in slot order
                rank += 1
            else:
                s_ = prim_base + off                                    # one primitive per leaf, contiguous in slot order
                off += 1
                assert s_ < n_bvh
                seen[s_] += 1
                plo, phi = slot_box(world, int(prim[s_]))
                # quantised boxes are each conservative for the primitives below them (a child's own
                # grid may stick out of its parent's slot box by a quantum, which is harmless)
                for blo, bhi in anc + [(lo, hi)]:
                    assert np.all(plo >= blo) and np.all(phi <= bhi), "a box on the path does not contain the primitive"
    assert np.all(seen[:n_bvh] == 1), "every BVH primitive must be in exactly one wide leaf"
    assert np.all(node_seen == 1), "every wide node must be reachable exactly once"


@pytest.mark.parametrize("n,ntri,ground", [(1, 0, False), (2, 0, False), (5, 3, False), (64, 16, True), (1000, 200, True), (30000, 0, False)])
@pytest.mark.parametrize("refit_mode", [0, 1, 2])
def test_wide_tree_invariants(pkg, synth, n, ntri, ground, refit_mode):
    d = scenes.random_scene(pkg, synth, n, n_tris=ntri, seed=n, with_walls=n >= 64, with_ground=ground)
    world = scenes.flatten_world(d)
    gs = scenes.GpuScene(pkg, d, options={"refit_mode": refit_mode})
    check_wide_tree(gs, world)
    gs.scene.build()  # rebuild over the same buffers: same invariants
    check_wide_tree(gs, world)
    gs.close()


@pytest.mark.parametrize("n,ntri,ground", [(1, 0, False), (5, 3, False), (1000, 200, True), (30000, 0, False)])
def test_wide_tree_group_collapse(pkg, synth, oracle, n, ntri, ground):
    """option collapse_mode = 1 (k_collapse8: eight lanes per wide node, all internal children opened per round):
    another valid 8-wide tree over the same primitives — same invariants, same image bit for bit."""
    d = scenes.random_scene(pkg, synth, n, n_tris=ntri, seed=n, with_walls=n >= 64, with_ground=ground)
    world = scenes.flatten_world(d)
    gs = scenes.GpuScene(pkg, d, options={"collapse_mode": 1})
    check_wide_tree(gs, world)
    cam = scenes.config_camera(pkg, "VK")
    g = gs.render(cam, 96, 64, 2, 5, 5)
    o = oracle.render(world, cam, 96, 64, 2, 5, 5, accel=1)
    assert np.array_equal(g["prim_id"], o["prim_id"]) and np.array_equal(g["accum"].view(np.uint32), o["accum"].view(np.uint32))
    gs.close()


def test_wide_tree_duplicate_keys(pkg):
    c = np.zeros((300, 3), np.float32)
    c[:200] = [0.1, 0.1, 0.1]
    c[200:] = np.random.default_rng(0).uniform(-0.4, 0.4, (100, 3))
    r = np.linspace(0.01, 0.03, 300).astype(np.float32)
    d = scenes.SceneDesc(scenes.CONFIG_ROUGHS, scenes.CONFIG_METALS)
    g = d.add_spheres(c, r)
    d.add_instance(g, None, 0, 0, 0)
    gs = scenes.GpuScene(pkg, d)
    check_wide_tree(gs, scenes.flatten_world(d))
    gs.close()


def test_error_paths(pkg):
    ctx = pkg.Context(0)
    sc = pkg.Scene(ctx)
    cam = pkg.camera_configure([1, 0, 0], [0, 0, 0], [0, 0, 1], False)
    with pytest.raises(pkg.RtError) as e:
        sc.render(cam, 8, 8, 1, 5, 0)
    assert e.value.code == pkg.RT_ERR_STATE
    with pytest.raises(pkg.RtError) as e:
        sc.add_instance(3)
    assert e.value.code == pkg.RT_ERR_INVALID
    import torch
    c = torch.zeros((4, 3), device="cuda"); r = torch.ones(4, device="cuda")
    g = sc.add_spheres(c.data_ptr(), r.data_ptr(), 4)
    sc.add_instance(g, pkg.construct_transform([0, 0, 0], [0, 0, 0], [1, 2, 3]))
    with pytest.raises(pkg.RtError) as e:
        sc.build()
    assert e.value.code == pkg.RT_ERR_UNSUPPORTED
    sc.close()
    ctx.close()
