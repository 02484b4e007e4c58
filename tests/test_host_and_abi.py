"""CPU tests of the host-side mirror of the reference (Scene / packers / BVH builder / Camera),
the byte layouts, and the C ABI surface: the library must load without a GPU and export every
symbol include/wgpt.h declares; with no CUDA device wgpt_create must FAIL (no CPU fallback)."""
import ctypes as C
import os
import re
import numpy as np
import pytest

from webgpu_path_tracer_b200 import Scene, Camera, Transform, ObjReader, api
from webgpu_path_tracer_b200 import glmatrix as glm
from webgpu_path_tracer_b200.bvh import build_flat_bvh, build_flat_bvh_recursive, triangle_boxes
from webgpu_path_tracer_b200 import meshgen

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_default_scene_inventory(default_scene):
    """19 spheres, 8 quads, 12 triangles, 23 BVH nodes, 31 materials, 28 transforms."""
    s, cam = default_scene
    b = s.buffers()
    assert b[1].size == 19 * 8 and b[2].size == 8 * 20 and b[5].size == 12 * 24
    assert b[9].size == 23 * 12 and b[8].size == 31 * 16 and b[7].size == 28 * 32 and b[6].size == 4
    assert b[6].dtype == np.int32 and all(b[k].dtype == np.float32 for k in (1, 2, 5, 7, 8, 9))
    d = s.material_dict
    assert (d["default"], d["red"], d["green"], d["blue"], d["white"], d["glossywhite"], d["black"], d["glass"]) == (0, 20, 21, 22, 23, 24, 25, 26)
    assert (d["tWall"], d["fWall"], d["dragonMat"], d["glassBox"]) == (27, 28, 29, 30)
    mats = b[8].reshape(-1, 16)
    assert np.isnan(mats[27, 13])                      # eta `undefined` -> NaN (scene.js:119)
    assert mats[27, 8] == 2.0                          # the light quad's emission
    sph = b[1].reshape(-1, 8)
    assert list(sph[:, 6]) == list(range(1, 20)) and list(sph[:, 4]) == list(range(19))
    quads = b[2].reshape(-1, 20)
    assert list(quads[:, 11]) == list(range(19, 27)) and list(quads[:, 19]) == [27, 25, 20, 21, 23, 24, 23, 28]
    assert list(b[6]) == [12, 0, 27, 30]


def test_struct_strides_match_header():
    """80/32/80/16/96/16/128/64/48 bytes."""
    assert api.STRIDES == {0: 80, 1: 32, 2: 80, 3: 16, 5: 96, 6: 16, 7: 128, 8: 64, 9: 48}
    hdr = open(os.path.join(ROOT, "include", "wgpt.h")).read()
    for name, num in (("UNIFORMS", 0), ("SPHERES", 1), ("QUADS", 2), ("FRAMEBUFFER", 3), ("TRIANGLES", 5),
                      ("MESHES", 6), ("TRANSFORMS", 7), ("MATERIALS", 8), ("BVH", 9)):
        assert re.search(r"#define WGPT_BINDING_%s\s+%d\b" % (name, num), hdr)


def test_quad_record():
    s = Scene(default=False)
    q = s.add_quad([-1, 1, -1], [3, 0, 0], [0, 0, 2], 7)
    d = np.asarray(q.data, np.float32)
    # n = u x v = (0,-6,0); normal = (0,-1,0); D = dot(normal,Q) = -1; w = n/dot(n,n) = (0,-1/6,0)
    assert list(d[12:15]) == [0, -1, 0] and d[15] == -1 and np.allclose(d[16:19], [0, -1 / 6, 0]) and d[19] == 7
    assert d[3] == -1 and d[7] == 0 and d[11] == 0


def test_transform_and_camera():
    t = Transform()
    t.update(t.scale(1.2, 1.2, 1.2), t.translate(0.1, -1.05, 0))
    m = np.asarray(t.modelMatrix).reshape(4, 4).T
    assert np.allclose(m @ [1, 1, 1, 1], [1.3, 0.15, 1.2, 1], atol=1e-6)          # scale first, then translate
    inv = np.asarray(t.invModelMatrix).reshape(4, 4).T
    assert np.allclose(inv @ m, np.eye(4), atol=1e-6)
    r = Transform()
    r.update(r.rotate(np.pi / 2, [0, 1, 0]))
    assert np.allclose(np.asarray(r.modelMatrix).reshape(4, 4).T @ [1, 0, 0, 0], [0, 0, -1, 0], atol=1e-6)
    cam = Camera()
    cam.set_camera([0.5, 0, 2.5], [0.5, 0, 0], [0, 1, 0])
    assert np.allclose(cam.viewMatrix, [1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0.5, 0, 2.5, 1])
    assert all(float(np.float32(x)) == x for x in cam.viewMatrix)


def test_obj_reader_rules():
    text = "# c\no X\nv 0 0 0\nv 1 0 0\nv 0 1 0\nvn 0 0 1\nvt 0 0\ns 0\nf 1/1/1 2/1/1 3/1/1\n  f 3/1/1 2/1/1 1/1/1  \n"
    m = ObjReader.parse(text)
    assert m["vertices"].dtype == np.float32 and m["vertices"].size == 18 and m["normals"].size == 18
    assert list(m["vertices"][:9]) == [0, 0, 0, 1, 0, 0, 0, 1, 0] and list(m["vertices"][9:12]) == [0, 1, 0]


def test_bvh_builder_matches_literal_recursion(golden_meshes):
    rng = np.random.default_rng(5)
    for name in ("cube", "cone", "hole", "monkey_968"):
        v = golden_meshes[name + "_vertices"].reshape(-1, 3, 3).astype(np.float64)
        mn, mx = triangle_boxes(v[:, 0], v[:, 1], v[:, 2])
        a, oa = build_flat_bvh(mn, mx)
        b, ob = build_flat_bvh_recursive(mn, mx)
        assert np.array_equal(a, b) and np.array_equal(oa, ob)
    # many equal keys (stability of the sort matters) and a non-power-of-two count
    pts = np.floor(rng.uniform(0, 4, (777, 3)))
    a, oa = build_flat_bvh(pts - 0.25, pts + 0.25)
    b, ob = build_flat_bvh_recursive(pts - 0.25, pts + 0.25)
    assert np.array_equal(a, b) and np.array_equal(oa, ob)


def test_bvh_node_invariants(golden_meshes):
    v = golden_meshes["monkey_968_vertices"].reshape(-1, 3, 3).astype(np.float64)
    mn, mx = triangle_boxes(v[:, 0], v[:, 1], v[:, 2])
    nodes, order = build_flat_bvh(mn, mx)
    n = v.shape[0]
    assert nodes.shape == (2 * n - 1, 12) and sorted(order) == list(range(n))
    leaf = nodes[:, 7] == 2
    assert leaf.sum() == n and (nodes[leaf, 9] == 1).all() and (nodes[leaf, 3] == -1).all()
    assert sorted(nodes[leaf, 8].astype(int)) == list(range(n))
    inner = np.nonzero(~leaf)[0]
    right = nodes[inner, 3].astype(int)
    assert (right > inner + 1).all() and (right < 2 * n - 1).all() and set(nodes[inner, 11].astype(int)) <= {0, 1, 2}
    # children boxes are inside the parent's; subtree [i, skip) holds (skip - i + 1)/2 triangles
    for i in inner[:200]:
        for c in (i + 1, int(nodes[i, 3])):
            assert (nodes[c, 0:3] >= nodes[i, 0:3]).all() and (nodes[c, 4:7] <= nodes[i, 4:7]).all()
        end = int(nodes[i, 10]) if nodes[i, 10] >= 0 else 2 * n - 1
        assert leaf[i:end].sum() == (end - i + 1) // 2
    thin = (mx - mn).min(axis=1) < 0.0001 / 2 - 1e-12          # pad() widens boxes thinner than 5e-5 (AABB.js:35-51)
    assert not thin.any()


def test_meshgen_counts():
    m = meshgen.noise_sphere(69451, 2)
    assert m["vertices"].size // 9 == 69432 and m["normals"].shape == m["vertices"].shape
    n = m["normals"].reshape(-1, 3)
    assert np.allclose(np.linalg.norm(n, axis=1), 1, atol=1e-4)
    m2 = meshgen.noise_sphere(69451, 2)
    assert np.array_equal(m["vertices"], m2["vertices"])


def test_library_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "wgpt.h")).read()
    declared = set(re.findall(r"\b(wgpt_[a-z_0-9]+)\s*\(", hdr))
    assert declared == set(api.EXPORTS), declared ^ set(api.EXPORTS)
    lib = api.load_library()
    for name in declared:
        assert hasattr(lib, name), name


def test_no_cpu_fallback_without_gpu():
    """On a box without a CUDA device the product path must fail loudly."""
    try:
        C.CDLL("libcuda.so.1")
        have_driver = True
    except OSError:
        have_driver = False
    if have_driver:
        pytest.skip("a CUDA driver is present; the no-GPU failure path is exercised in the build container")
    with pytest.raises(api.WgptError, match="no CUDA device|CUDA"):
        api.PathTracer(16, 16)


def test_config_validation_runs_before_cuda():
    lib = api.load_library()
    cfg = api.Config()
    lib.wgpt_default_config(C.byref(cfg))
    assert cfg.struct_size == C.sizeof(api.Config) and cfg.num_samples == 1 and cfg.max_bounces == 100 and cfg.stack_size == 20
    h = C.c_void_p()
    cfg.width = 0
    assert lib.wgpt_create(C.byref(cfg), C.byref(h)) == -1 and b"width" in lib.wgpt_last_error(None)
    cfg.width, cfg.height = 5000, 5000                  # > 2^24 pixels: f32(pixelIndex) would not be exact
    assert lib.wgpt_create(C.byref(cfg), C.byref(h)) == -1
    cfg.width = cfg.height = 64
    cfg.stack_size = 65
    assert lib.wgpt_create(C.byref(cfg), C.byref(h)) == -1


def test_product_package_never_imports_oracle():
    pkg = os.path.join(ROOT, "webgpu_path_tracer_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert "oracle/" not in src.replace("oracle/ovec.h", "").replace("oracle/)", "") or f in ("api.py", "__init__.py", "pvec.cuh"), f
                assert "import oracle" not in src and "from oracle" not in src, f


def test_png_and_ppm_writers_round_trip(tmp_path):
    """N4: the canvas is a file here; the PNG the host layer writes decodes back to the same pixels."""
    from webgpu_path_tracer_b200 import imageio
    rng = np.random.default_rng(3)
    img = rng.random((37, 53, 4)).astype(np.float32)
    want = np.clip(np.rint(img * 255.0), 0, 255).astype(np.uint8)
    for alpha in (False, True):
        got = imageio.decode_png(imageio.encode_png(img, alpha=alpha))
        assert np.array_equal(got, want if alpha else want[:, :, :3])
    p = tmp_path / "x.ppm"
    imageio.write_ppm(p, img)
    raw = p.read_bytes()
    assert raw.startswith(b"P6\n53 37\n255\n") and raw[len(b"P6\n53 37\n255\n"):] == want[:, :, :3].tobytes()


def test_napi_addon_source_passes_a_syntax_check_and_binds_only_declared_entry_points():
    """N3: the Node addon is source only here (no node, no node_api.h).  It must at least be valid
    C++ against the N-API signatures it uses (stand-in header under tests/napi_stub) and call nothing that
    include/wgpt.h does not declare."""
    import re, shutil, subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    src = os.path.join(root, "bindings", "napi", "addon.cc")
    if shutil.which("g++"):
        subprocess.check_call(["g++", "-std=c++17", "-fsyntax-only", "-Wall", "-Werror", "-I" + os.path.join(root, "tests", "napi_stub"),
                               "-I" + os.path.join(root, "include"), src])
    header = open(os.path.join(root, "include", "wgpt.h")).read()
    used = set(re.findall(r"\b(wgpt_[a-z_0-9]+)\s*\(", open(src).read()))
    assert used and all(re.search(r"\b%s\s*\(" % u, header) for u in used), used
    shim = open(os.path.join(root, "bindings", "napi", "wgpt-utils.js")).read()
    for method in ("createUniformBuffer", "createStorageBuffer_WriteOnly", "createStorageBuffer_ReadWrite", "createVertexBuffer",
                   "createShaderModule", "createComputePipeline", "createRenderPipeline", "createRenderPassDescriptor",
                   "computePass", "renderPass", "createBindGroup", "writeBuffer", "init"):
        assert re.search(r"\b%s\s*\(" % method, shim), method        # the call shapes renderer.js uses


def test_pfm_round_trip(tmp_path):
    from webgpu_path_tracer_b200 import imageio
    rng = np.random.default_rng(0)
    img = rng.uniform(0, 30, (5, 7, 4)).astype(np.float32)
    img[0, 0, 0] = np.inf
    p = tmp_path / "x.pfm"
    imageio.write_pfm(p, img)
    assert p.read_bytes().startswith(b"PF\n7 5\n-1.0\n")
    assert np.array_equal(imageio.read_pfm(p), img[:, :, :3])


def test_shard_grid_assignment_covers_the_job_exactly_once():
    """tile x strata grid (sharding.shard_of_rank, the rule wgpt_group_create uses): over all ranks every (pixel, frame)
    of a job is rendered exactly once, for every split bench.py uses."""
    from webgpu_path_tracer_b200 import sharding
    W, H, first, n = 40, 37, 3, 11
    for world in (1, 2, 4, 8):
        for tiles in sorted({1, sharding.default_tile_groups(world), world}):
            count = np.zeros((n, H * W), np.int32)
            for rank in range(world):
                tr, tw, sr, sw = sharding.shard_of_rank(rank, world, tiles)
                assert (tw, sw) == (tiles, world // tiles)
                mask = sharding.tile_mask(W, H, tr, tw, 8)
                for f in sharding.strata_frames(first, n, sr, sw):
                    count[f - first] += mask
            assert (count == 1).all(), (world, tiles)
    assert [sharding.default_tile_groups(w) for w in (1, 2, 4, 8)] == [1, 1, 2, 4]
    with pytest.raises(ValueError):
        sharding.shard_of_rank(0, 6, 4)
