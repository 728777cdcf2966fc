"""CPU tests that run the reference itself (build container only: /root/reference; deselected on the GPU box where the
tree does not exist).  They tie two host-side pieces of the product to the reference's own objects rather than to
stand-ins: world.flatten_frame on the data.Frame objects chunk_update built, and the C++ Goxel loader on the three shipped
models."""
import os

import numpy as np
import pytest

from oracle import refharness as rh

pytestmark = [pytest.mark.reference, pytest.mark.skipif(not rh.available(), reason="reference tree not present")]


@pytest.fixture(scope="module")
def ref_world():
    ref = rh.load_reference()
    rh.restore_settings()
    rh.override_settings(culling=False, dist_move=0)
    cam = ref.Camera()
    cam.pos, cam.rot = ref.data.player.cam_pos, ref.data.player.cam_rot
    st = rh.build_chunks(cam)
    yield ref, cam, st
    rh.restore_settings()


def test_flatten_frame_on_the_references_own_frames(ref_world):
    """the data.Frame objects the reference's chunk_update produced (greedy data6 boxes + data3
    points, three LODs per chunk) -> world.flatten_frame == the brute-force grid refharness.extract_world reads cell by
    cell through Frame.get_voxel; for the LOD frames every cell's lookup is compared (key = floor(p) // r)."""
    from python_raytracer_b200.world import flatten_frame
    ref, cam, st = ref_world
    cs = ref.data.settings.chunk_size
    origin, cells, chunk_res, table, names = rh.extract_world(st, cam)
    ids = {}

    def mat_id(m):                                       # same numbering as extract_world: by the optical fields
        row = (m.albedo.r, m.albedo.g, m.albedo.b, m.roughness, m.absorption, m.ior, m.energy)
        if id(m) not in ids:
            hit = [k for k in range(1, len(table)) if tuple(table[k]) == row]
            assert hit, row
            ids[id(m)] = hit[0]
        return ids[id(m)]

    posts = [p for p, v in st.chunks.items() if v]
    boxes = sum(len(st.chunks[p][0].data6) for p in posts)
    assert len(posts) >= 150 and boxes >= 1000                          # 198 chunks, 1 415 greedy boxes when surveyed
    for post in posts:
        flat = flatten_frame(st.chunks[post][0], post, cs, mat_id)
        lo = np.array(post) - origin
        want = cells[lo[0]:lo[0] + cs, lo[1]:lo[1] + cs, lo[2]:lo[2] + cs]
        # material ids may be numbered differently only if two materials share every field; compare through the table
        assert np.array_equal(table[flat], table[want]), post
    vec3 = ref.lib.vec3
    rng = np.random.default_rng(3)
    for post in [posts[i] for i in rng.choice(len(posts), 24, replace=False)]:
        for lod in (1, 2):
            frame = st.chunks[post][lod]
            r = frame.resolution
            flat = flatten_frame(frame, post, cs, mat_id)
            for x in range(post[0], post[0] + cs):
                for y in range(post[1], post[1] + cs):
                    for z in range(post[2], post[2] + cs):
                        m = frame.get_voxel(vec3(x, y, z))
                        q = ((x // r) * r - post[0], (y // r) * r - post[1], (z // r) * r - post[2])
                        got = int(flat[q]) if min(q) >= 0 else 0        # the device's lookup
                        assert got == (mat_id(m) if m is not None else 0), (post, lod, x, y, z)


def test_goxel_loader_on_the_three_shipped_models(ref_world):
    """parity with Sprite.load (data.py:271-289) on mods/default/voxels/*.txt.gz (castle: 209 434 lines)."""
    import sys
    from python_raytracer_b200.goxel import load_goxel
    ref, cam, st = ref_world
    mod = sys.modules["mods.default.init"]
    cases = [("castle.txt.gz", "castle_spr", {"000000": "mat_metal", "3f3f3f": "mat_stone_dark", "7f7f7f": "mat_stone_gray",
                                                "bfbfbf": "mat_stone_light", "ffffff": "mat_stone_marble"}),
             ("material.txt.gz", "material_glass_spr", {"7f7f7f": "mat_material", "ffffff": "mat_material_glass"}),
             ("player.txt.gz", "player_spr", {"7f7f7f": "mat_player"})]
    for fname, spr_name, keymap in cases:
        spr = getattr(mod, spr_name)
        want = spr.frames[0].get_voxels()                                # what Sprite.load stored (resolution 1)
        keys = list(keymap)
        xyz, m = load_goxel(os.path.join(rh.REFERENCE_ROOT, "mods", "default", "voxels", fname), int(spr.size.x), keys)
        got = {}
        for p, k in zip(xyz.tolist(), m.tolist()):                      # file order, later rows override (dict semantics)
            got[tuple(p)] = getattr(mod, keymap[keys[k - 1]])
        assert len(got) == len(want) > 0, fname
        assert all(want.get(p) is mat for p, mat in got.items()), fname
