"""TEST INFRASTRUCTURE: a SECOND, independent restatement of the reference's per-pixel path, in numpy.

Written from the Rust sources (file:line below), not from oracle/rt_oracle.cpp, and in a different style: every
pixel of the frame at once in float32 arrays, the BVH walked as a recursive descent over the uploaded nodes with a
mask of the rays that still pass every ancestor's slab test (the reference never prunes by distance, so the set of
nodes a ray visits is exactly "all ancestors and the node pass ray_aabb_intersect", shader/src/bvh.rs:40-85; the
64-entry stack quirk cannot trigger on the small trees used here).  Two restatements that agree bit for bit make a
transcription error in either less likely; neither pins the arithmetic of the third-party crates (glam, spirv-std),
whose published scalar semantics both use (oracle/rt_oracle_math.h).

Only tests/ imports this.  numpy float32 arithmetic is IEEE single (no contraction); libm's tanf is the same
function the oracle calls (the shader's tan is the GPU's)."""
import ctypes

import numpy as np

F = np.float32
_libm = ctypes.CDLL("libm.so.6")
_libm.tanf.restype = ctypes.c_float
_libm.tanf.argtypes = [ctypes.c_float]

MIN_RAY_DISTANCE = F(1e-5)      # shared/src/lib.rs:27
F32_MAX = np.finfo(np.float32).max
INVALID = 0xFFFFFFFF


# ---- glam scalar-math vector semantics (tuples of float32 arrays) ---------------------------------------------------
def dot(a, b):
    return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]


def cross(a, b):
    return (a[1] * b[2] - b[1] * a[2], a[2] * b[0] - b[2] * a[0], a[0] * b[1] - b[0] * a[1])


def length(a):
    return np.sqrt(dot(a, a))


def normalize(a):
    with np.errstate(all="ignore"):
        r = F(1.0) / length(a)
        return (a[0] * r, a[1] * r, a[2] * r)


def add(a, b):
    return (a[0] + b[0], a[1] + b[1], a[2] + b[2])


def sub(a, b):
    return (a[0] - b[0], a[1] - b[1], a[2] - b[2])


def scale(a, s):
    return (a[0] * s, a[1] * s, a[2] * s)


def vec(x):
    return (F(x[0]), F(x[1]), F(x[2]))


def f16_to_f32(bits16):
    return np.asarray(bits16, np.uint16).view(np.float16).astype(np.float32)


class Frame:
    """The bound buffers of one dispatch, read the way SceneAccessor reads them (shader/src/scene_access.rs)."""

    def __init__(self, meta_u32, offsets, triangles, materials_bound, camera, width, height, bound_len):
        self.u = np.asarray(meta_u32, np.uint32)
        self.f = self.u.view(np.float32)
        self.off = offsets
        self.tris = triangles
        self.mats = materials_bound
        self.cam = camera
        self.w, self.h = int(width), int(height)
        self.bound_len = int(bound_len)

    def o(self, name):
        return int(self.off[name])


# ---- shader/src/ray.rs:22-53 (legacy) and shader/src/wavefront.rs:75-112 (wavefront mode) ----------------------------
def camera_rays(fr, wavefront):
    ys, xs = np.mgrid[0:fr.h, 0:fr.w]
    px, py = xs.astype(np.float32).ravel(), ys.astype(np.float32).ravel()
    res0, res1 = F(fr.w), F(fr.h)  # resolution as f32; `as u32` and back is the identity for these sizes
    u, v = (px + F(0.5)) / res0, (py + F(0.5)) / res1
    aspect = res0 / res1
    fov_scale = F(_libm.tanf(F(fr.cam["fov"]) * F(0.5) * F(np.pi) / F(180.0)))
    cx = (u * F(2.0) - F(1.0)) * aspect * fov_scale
    cy = (F(1.0) - v * F(2.0)) * fov_scale
    fwd, up = vec(fr.cam["direction"]), vec(fr.cam["up"])
    right = cross(fwd, up)
    true_up = cross(right, fwd)
    d = add(add(fwd, scale(right, cx)), scale(true_up, cy))
    d = normalize(d)
    if not wavefront:
        d = normalize(d)  # Ray::new normalises again (ray.rs:14-19, 52)
    o = tuple(np.full_like(px, c) for c in vec(fr.cam["position"]))
    return o, d


# ---- shader/src/intersection.rs:52-87 ---------------------------------------------------------------------------------
def sphere_pass(fr, o, d, closest):
    n = o[0].shape[0]
    hit = np.zeros(n, bool)
    t_best = np.full(n, F32_MAX, np.float32)
    idx = np.full(n, INVALID, np.uint32)
    base = fr.o("spheres_offset")
    for i in range(fr.o("spheres_count")):
        w = base + i * 5
        c, r = (fr.f[w], fr.f[w + 1], fr.f[w + 2]), fr.f[w + 3]
        oc = sub(o, c)
        a = dot(d, d)
        b = F(2.0) * dot(oc, d)
        cc = dot(oc, oc) - r * r
        disc = b * b - F(4.0) * a * cc
        with np.errstate(all="ignore"):
            sq = np.sqrt(disc)
            t1 = (-b - sq) / (F(2.0) * a)
            t2 = (-b + sq) / (F(2.0) * a)
        t = np.where(t1 > MIN_RAY_DISTANCE, t1, t2)
        ok = (disc >= 0) & (t > MIN_RAY_DISTANCE) & (t < closest)
        closest = np.where(ok, t, closest)
        t_best = np.where(ok, t, t_best)
        idx = np.where(ok, np.uint32(i), idx)
        hit |= ok
    return hit, t_best, idx, closest


# ---- shader/src/intersection.rs:151-164 ------------------------------------------------------------------------------
def slab(o, d, bmin, bmax):
    with np.errstate(all="ignore"):
        inv = (F(1.0) / d[0], F(1.0) / d[1], F(1.0) / d[2])
        t1 = tuple((bmin[k] - o[k]) * inv[k] for k in range(3))
        t2 = tuple((bmax[k] - o[k]) * inv[k] for k in range(3))
        tmin = tuple(np.fmin(t1[k], t2[k]) for k in range(3))
        tmax = tuple(np.fmax(t1[k], t2[k]) for k in range(3))
        tmin_max = np.fmax(np.fmax(tmin[0], tmin[1]), tmin[2])
        tmax_min = np.fmin(np.fmin(tmax[0], tmax[1]), tmax[2])
        return (tmax_min >= 0) & (tmin_max <= tmax_min)


# ---- shader/src/triangle_access.rs:18-61 + intersection.rs:91-138 ----------------------------------------------------
def triangle_test(fr, tri_index, tpb, o, d, closest, active):
    """Returns (accepted mask, t, material id, normal) of one triangle for the active rays.  `tpb` is the push
    constants' triangles_per_buffer (the shader's divisor); fr.bound_len is the length of every bound buffer."""
    buf, local = tri_index // tpb, tri_index % tpb
    n_buf = max(1, (len(fr.tris) + fr.bound_len - 1) // fr.bound_len)
    if buf > 2 or local >= fr.bound_len:
        return None
    # binding k is buffer k if it exists, else buffer 0 (src/renderer.rs:680-693); every buffer holds exactly
    # bound_len zero-initialised entries (src/buffers.rs:93-99)
    flat = (buf if buf < n_buf else 0) * fr.bound_len + local
    if flat < len(fr.tris):
        tri = fr.tris[flat]
        vi, mat = (int(tri["v0_index"]), int(tri["v1_index"]), int(tri["v2_index"])), int(tri["material_id"])
    else:
        vi, mat = (0, 0, 0), 0  # the zero-filled tail of a bound buffer
    vb = fr.o("vertices_offset")
    v0, v1, v2 = [(fr.f[vb + k * 3], fr.f[vb + k * 3 + 1], fr.f[vb + k * 3 + 2]) for k in vi]
    e1, e2 = sub(v1, v0), sub(v2, v0)
    h = cross(d, e2)
    a = dot(e1, h)
    with np.errstate(all="ignore"):
        ok = active & ~(np.abs(a) < MIN_RAY_DISTANCE)
        f = F(1.0) / a
        s = sub(o, v0)
        u = f * dot(s, h)
        ok &= ~((u < 0) | (u > F(1.0)))
        q = cross(s, e1)
        v = f * dot(d, q)
        ok &= ~((v < 0) | (u + v > F(1.0)))
        t = f * dot(e2, q)
        ok &= (t > MIN_RAY_DISTANCE) & (t < closest)
    nrm = normalize(cross(e1, e2))
    return ok, t, mat, nrm


class TriResult:
    def __init__(self, n):
        self.hit = np.zeros(n, bool)
        self.t = np.full(n, F32_MAX, np.float32)
        self.mat = np.full(n, INVALID, np.uint32)
        self.prim = np.full(n, INVALID, np.uint32)
        self.nrm = [np.zeros(n, np.float32) for _ in range(3)]

    def take(self, ok, t, mat, prim, nrm):
        self.hit |= ok
        self.t = np.where(ok, t, self.t)
        self.mat = np.where(ok, np.uint32(mat), self.mat)
        self.prim = np.where(ok, np.uint32(prim), self.prim)
        for k in range(3):
            self.nrm[k] = np.where(ok, nrm[k], self.nrm[k])


# ---- shader/src/bvh.rs:18-133 as a recursive descent with ray masks ----------------------------------------------------
def bvh_pass(fr, tpb, o, d, closest):
    res = TriResult(o[0].shape[0])
    nb, n_nodes, n_idx = fr.o("bvh_nodes_offset"), fr.o("bvh_nodes_count"), fr.o("triangle_indices_count")
    ib = fr.o("triangle_indices_offset")
    state = {"closest": closest}

    def visit(node, mask, depth):
        assert depth < 60, "tree too deep for this restatement (the reference's stack quirk would matter)"
        if node == INVALID or node >= n_nodes or not mask.any():
            return
        w = nb + node * 12
        mask = mask & slab(o, d, (fr.f[w], fr.f[w + 1], fr.f[w + 2]), (fr.f[w + 4], fr.f[w + 5], fr.f[w + 6]))
        if not mask.any():
            return
        left, right, start, count = (int(fr.u[w + 8 + k]) for k in range(4))
        if left == INVALID:  # leaf (bvh.rs:60)
            # test_leaf_triangles keeps its own running minimum seeded with the traversal's, and the leaf's result
            # replaces the traversal's only if it hit: same as updating one running minimum
            for i in range(count):
                if start + i >= n_idx:
                    break
                tri_index = int(fr.u[ib + start + i])
                r = triangle_test(fr, tri_index, tpb, o, d, state["closest"], mask)
                if r is None:
                    continue
                ok, t, mat, nrm = r
                state["closest"] = np.where(ok, t, state["closest"])
                res.take(ok, t, mat, tri_index, nrm)
        else:  # right is pushed first, so left is popped first (bvh.rs:74-83)
            visit(left, mask, depth + 1)
            visit(right, mask, depth + 1)

    visit(0, np.ones(o[0].shape[0], bool), 0)
    return res


# ---- shader/src/lib.rs:272-296 -----------------------------------------------------------------------------------------
def brute_force_pass(fr, tpb, triangle_count, o, d, closest):
    res = TriResult(o[0].shape[0])
    everyone = np.ones(o[0].shape[0], bool)
    for i in range(triangle_count):
        r = triangle_test(fr, i, tpb, o, d, closest, everyone)
        if r is None:
            continue
        ok, t, mat, nrm = r
        closest = np.where(ok, t, closest)
        res.take(ok, t, mat, i, nrm)
    return res


# ---- shader/src/material.rs + lighting.rs + lib.rs:300-338 -------------------------------------------------------------
def shade(fr, point, normal, material_id, channel):
    n = point[0].shape[0]
    n_bound = len(fr.mats)
    valid = material_id < n_bound
    mid = np.where(valid, material_id, 0).astype(np.int64)
    m = fr.mats[mid]
    albedo = tuple(m["albedo"][:, k].astype(np.float32) for k in range(3))
    emission = tuple(m["emission"][:, k].astype(np.float32) for k in range(3))
    mr, it = m["metallic_roughness_f16"].astype(np.uint32), m["ior_transmission_f16"].astype(np.uint32)
    metallic = f16_to_f32(mr & 0xFFFF)
    ior, transmission = f16_to_f32(it & 0xFFFF), f16_to_f32(it >> 16)
    total = scale(albedo, F(0.1))
    total = add((F(0.0), F(0.0), F(0.0)), total)
    lb = fr.o("lights_offset")
    with np.errstate(all="ignore"):
        for li in range(fr.o("lights_count")):
            w = lb + li * 13
            lpos = (fr.f[w], fr.f[w + 1], fr.f[w + 2])
            ltype = int(fr.u[w + 3])
            lcol = (fr.f[w + 4], fr.f[w + 5], fr.f[w + 6])
            lint = fr.f[w + 7]
            ldir = tuple(np.full(1, fr.f[w + 8 + k], np.float32) for k in range(3))
            nld = tuple(-c for c in normalize(ldir))
            dir_i = np.fmax(dot(normal, nld), F(0.0)) * lint
            to_light = sub(lpos, point)
            dist = length(to_light)
            pdir = normalize(to_light)
            att = F(1.0) / (F(1.0) + dist * dist * F(0.01))
            att = att.astype(np.float16).astype(np.float32)  # f32 -> f16 -> f32 (lighting.rs:125-127)
            point_i = np.fmax(dot(normal, pdir), F(0.0)) * lint * att
            spot_i = point_i * np.fmax(dot(nld, pdir), F(0.0))
            final = dir_i * F(ltype == 0) + point_i * F(ltype == 1) + spot_i * F(ltype == 2)
            diffuse = tuple(c / F(np.pi) for c in albedo)
            is_m = (metallic > F(0.5)).astype(np.float32)
            metal = scale(scale(albedo, final), F(0.5))
            diel = scale(diffuse, final)
            brdf = add(scale(metal, is_m), scale(diel, F(1.0) - is_m))
            ok = (final > 0).astype(np.float32)
            total = add(total, tuple(brdf[k] * lcol[k] * ok for k in range(3)))
        lighting = add(total, emission)
        T = np.fmin(np.fmax(transmission, F(0.0)), F(1.0))
        table = (F(-0.018), F(0.0), F(0.035), F(0.0))
        ior_c = ior + table[channel if channel < 3 else 3]
        disp = (ior_c - F(1.0)) / (ior - F(1.0))
        transmitted = (F(0.2) * disp, F(0.2) * disp, F(0.3) * disp)
        glass = add(scale(lighting, F(1.0) - T), scale(transmitted, T))
    out = tuple(np.where(T > 0, glass[k], lighting[k]) for k in range(3))
    magenta = (F(1.0), F(0.0), F(1.0))
    return tuple(np.where(valid, out[k], magenta[k]).astype(np.float32) for k in range(3))


# ---- shader/src/lib.rs:25-89, 92-149, 174-249 --------------------------------------------------------------------------
def render(bound, width, height, camera=None, wavefront=None, tpb=None, triangle_count=None):
    """All pixels of a frame, the three colour-channel passes.  wavefront = None (legacy) or (current, max) bounce
    depths.  Returns dict(kind, prim, t, material_id [H, W]; radiance [H, W, 3] = component c of pass c)."""
    sc = bound.scene
    fr = Frame(bound.meta, {k: bound.offsets[k] for k in bound.offsets.dtype.names} if hasattr(bound.offsets, "dtype") and bound.offsets.dtype.names
               else bound.offsets, np.ascontiguousarray(sc.triangles), bound.materials,
               sc.camera if camera is None else camera, width, height, bound.tpb)
    tpb = bound.tpb if tpb is None else tpb
    n = width * height
    black = wavefront is not None and wavefront[0] > wavefront[1]  # no iteration of lib.rs:116-145 runs
    o, d = camera_rays(fr, wavefront is not None)
    closest = np.full(n, F32_MAX - F(2.0), np.float32)
    s_hit, s_t, s_idx, closest = sphere_pass(fr, o, d, closest)
    if fr.o("bvh_nodes_count") > 0:
        tri = bvh_pass(fr, tpb, o, d, closest)
    else:
        tri = brute_force_pass(fr, tpb, len(sc.triangles) if triangle_count is None else triangle_count, o, d, closest)
    sphere_closer = s_t < tri.t
    use_sphere = (tri.hit & s_hit & sphere_closer) | (s_hit & ~tri.hit)
    use_tri = (tri.hit & s_hit & ~sphere_closer) | (tri.hit & ~s_hit)
    t = np.where(use_sphere, s_t, np.where(use_tri, tri.t, F32_MAX)).astype(np.float32)
    point = tuple(o[k] + d[k] * t for k in range(3))
    sb = fr.o("spheres_offset")
    si = np.where(s_hit, s_idx, 0).astype(np.int64)
    centre = tuple(fr.f[sb + si * 5 + k] for k in range(3))
    s_nrm = normalize(sub(point, centre))
    s_mat = fr.u[sb + si * 5 + 4]
    normal = tuple(np.where(use_sphere, s_nrm[k], tri.nrm[k]).astype(np.float32) for k in range(3))
    mat = np.where(use_sphere, s_mat, tri.mat).astype(np.uint32)
    hit = use_sphere | use_tri
    radiance = np.zeros((n, 3), np.float32)
    for ch in range(3):
        col = shade(fr, point, normal, np.where(hit, mat, 0).astype(np.uint32), ch)[ch]
        miss = F((0.1, 0.2, 0.3)[ch]) * F(1.0) if wavefront is not None else F(0.0)
        # wavefront mode multiplies by the camera ray's throughput (1, 1, 1)
        c = np.where(hit, col * F(1.0) if wavefront is not None else col, miss)
        radiance[:, ch] = F(0.0) if black else c
    kind = np.where(use_sphere, 1, np.where(use_tri, 2, 0)).astype(np.uint32)
    prim = np.where(use_sphere, s_idx, np.where(use_tri, tri.prim, INVALID)).astype(np.uint32)
    shp = (height, width)
    return dict(kind=kind.reshape(shp), prim=prim.reshape(shp), t=t.reshape(shp),
                material_id=np.where(hit, mat, INVALID).astype(np.uint32).reshape(shp), radiance=radiance.reshape(height, width, 3))


# ---- EXTENSION semantics (SPEC.md), restated from the SPEC text: per-sample paths with bounces ------------------------
def closest_hit(fr, tpb, o, d):
    """find_closest_intersection (lib.rs:174-249) for arbitrary rays -> (hit, t, point, normal, material id)."""
    n = o[0].shape[0]
    closest = np.full(n, F32_MAX - F(2.0), np.float32)
    s_hit, s_t, s_idx, closest = sphere_pass(fr, o, d, closest)
    if fr.o("bvh_nodes_count") > 0:
        tri = bvh_pass(fr, tpb, o, d, closest)
    else:
        tri = brute_force_pass(fr, tpb, len(fr.tris), o, d, closest)
    sphere_closer = s_t < tri.t
    use_sphere = (tri.hit & s_hit & sphere_closer) | (s_hit & ~tri.hit)
    use_tri = (tri.hit & s_hit & ~sphere_closer) | (tri.hit & ~s_hit)
    t = np.where(use_sphere, s_t, np.where(use_tri, tri.t, F32_MAX)).astype(np.float32)
    with np.errstate(all="ignore"):
        point = tuple(o[k] + d[k] * t for k in range(3))
        sb = fr.o("spheres_offset")
        si = np.where(s_hit, s_idx, 0).astype(np.int64)
        if fr.o("spheres_count"):
            centre = tuple(fr.f[sb + si * 5 + k] for k in range(3))
            s_nrm = normalize(sub(point, centre))
            s_mat = fr.u[sb + si * 5 + 4]
        else:
            s_nrm, s_mat = tri.nrm, tri.mat
    normal = tuple(np.where(use_sphere, s_nrm[k], tri.nrm[k]).astype(np.float32) for k in range(3))
    mat = np.where(use_sphere, s_mat, tri.mat).astype(np.uint32)
    return use_sphere | use_tri, t, point, normal, mat


class Rng:
    """SimpleRng (wavefront.rs:46-72) for an array of independent streams; `who` selects the streams that draw."""

    def __init__(self, seed):
        self.s = seed.astype(np.uint32)

    def next_f32(self, who):
        nxt = (self.s.astype(np.uint64) * np.uint64(1664525) + np.uint64(1013904223)).astype(np.uint32)  # wrapping
        self.s = np.where(who, nxt, self.s)
        return (self.s >> np.uint32(8)).astype(np.float32) / F(16777216.0)

    def next_f32_signed(self, who):
        return self.next_f32(who) * F(2.0) - F(1.0)


def render_path(bound, width, height, spp, max_depth, base_seed, camera=None):
    """SPEC.md steps 1-4 for every pixel.  Returns dict(accum [H, W, 4] float32, segments)."""
    sc = bound.scene
    fr = Frame(bound.meta, {k: bound.offsets[k] for k in bound.offsets.dtype.names}, np.ascontiguousarray(sc.triangles),
               bound.materials, sc.camera if camera is None else camera, width, height, bound.tpb)
    n = width * height
    ys, xs = np.mgrid[0:height, 0:width]
    xs, ys = xs.ravel(), ys.ravel()
    accum = np.zeros((n, 4), np.float32)
    segments = 0
    everyone = np.ones(n, bool)
    res0, res1 = F(width), F(height)
    fov_scale = F(_libm.tanf(F(fr.cam["fov"]) * F(0.5) * F(np.pi) / F(180.0)))
    fwd, up = vec(fr.cam["direction"]), vec(fr.cam["up"])
    right = cross(fwd, up)
    true_up = cross(right, fwd)
    with np.errstate(all="ignore"):
        for s in range(spp):
            seed = (np.uint64((base_seed + s) & 0xFFFFFFFF) + xs.astype(np.uint64) + ys.astype(np.uint64) * np.uint64(width)).astype(np.uint32)
            rng = Rng(seed)  # step 1
            if s == 0:       # step 2
                jx = jy = np.full(n, F(0.5), np.float32)
            else:
                jx = rng.next_f32(everyone)
                jy = rng.next_f32(everyone)
            u, v = (xs.astype(np.float32) + jx) / res0, (ys.astype(np.float32) + jy) / res1
            cx = (u * F(2.0) - F(1.0)) * (res0 / res1) * fov_scale
            cy = (F(1.0) - v * F(2.0)) * fov_scale
            d = normalize(add(add(fwd, scale(right, cx)), scale(true_up, cy)))
            o = tuple(np.full(n, c, np.float32) for c in vec(fr.cam["position"]))
            tp = [np.ones(n, np.float32) for _ in range(3)]
            L = [np.zeros(n, np.float32) for _ in range(3)]
            alive = everyone.copy()
            depth = 0
            while alive.any():
                segments += int(alive.sum())
                hit, t, point, normal, mat = closest_hit(fr, bound.tpb, o, d)   # 3.1
                miss = alive & ~hit
                sky = (F(0.1), F(0.2), F(0.3))
                for k in range(3):                                                # 3.2
                    L[k] = np.where(miss, L[k] + sky[k] * tp[k], L[k])
                alive &= hit
                valid = mat < len(fr.mats)
                mid = np.where(alive & valid, mat, 0).astype(np.uint32)
                C = [shade(fr, point, normal, mid, ch)[ch] for ch in range(3)]    # 3.3
                magenta = (F(1.0), F(0.0), F(1.0))
                for k in range(3):
                    c = np.where(valid, C[k], magenta[k])
                    L[k] = np.where(alive, L[k] + c * tp[k], L[k])
                alive &= valid
                if depth == max_depth:                                            # 3.4
                    break
                if depth >= 3:                                                    # 3.5
                    p = np.fmin(np.fmax(np.fmax(np.fmax(tp[0], tp[1]), tp[2]), F(0.05)), F(1.0))
                    r = rng.next_f32(alive)
                    alive &= ~(r > p)
                    for k in range(3):
                        tp[k] = np.where(alive, tp[k] / p, tp[k])
                r0 = rng.next_f32(alive)                                          # 3.6
                uvec = [np.zeros(n, np.float32) for _ in range(3)]
                need = alive.copy()
                for _ in range(8):
                    a = rng.next_f32_signed(need)
                    b = rng.next_f32_signed(need)
                    c = rng.next_f32_signed(need)
                    q = a * a + b * b + c * c
                    ok = need & (q > F(1e-6)) & (q <= F(1.0))
                    inv = F(1.0) / np.sqrt(q)
                    for k, comp in enumerate((a, b, c)):
                        uvec[k] = np.where(ok, comp * inv, uvec[k])
                    need &= ~ok
                cosi = dot(normal, d)                                              # 3.7
                nf = tuple(np.where(cosi < 0, normal[k], -normal[k]) for k in range(3))
                w = add(nf, uvec)
                small = dot(w, w) < F(1e-8)
                w = tuple(np.where(small, nf[k], w[k]) for k in range(3))
                w = normalize(w)
                m = fr.mats[mid.astype(np.int64)]                                  # 3.8
                albedo = tuple(m["albedo"][:, k].astype(np.float32) for k in range(3))
                mr, it = m["metallic_roughness_f16"].astype(np.uint32), m["ior_transmission_f16"].astype(np.uint32)
                metallic, roughness = f16_to_f32(mr & 0xFFFF), f16_to_f32(mr >> 16)
                ior, transmission = f16_to_f32(it & 0xFFFF), f16_to_f32(it >> 16)
                T = np.fmin(np.fmax(transmission, F(0.0)), F(1.0))
                refr = (T > 0) & (r0 < T)
                eta = np.where(cosi < 0, F(1.0) / ior, ior)
                c1 = -dot(d, nf)
                kk = F(1.0) - eta * eta * (F(1.0) - c1 * c1)
                two_dn = F(2.0) * dot(d, nf)
                mirror = tuple(d[k] - nf[k] * two_dn for k in range(3))
                bend = tuple(d[k] * eta + nf[k] * (eta * c1 - np.sqrt(kk)) for k in range(3))
                tir = kk < 0
                d_refr = normalize(tuple(np.where(tir, mirror[k], bend[k]) for k in range(3)))
                side_refr = np.where(tir, F(1.0), F(-1.0))
                gl = tuple(mirror[k] * (F(1.0) - roughness) + w[k] * roughness for k in range(3))
                bad = (dot(gl, gl) < F(1e-8)) | (dot(gl, nf) <= 0)
                d_metal = normalize(tuple(np.where(bad, w[k], gl[k]) for k in range(3)))
                is_metal = metallic > F(0.5)
                new_d = tuple(np.where(refr, d_refr[k], np.where(is_metal, d_metal[k], w[k])) for k in range(3))
                side = np.where(refr, side_refr, F(1.0))
                new_o = tuple(point[k] + nf[k] * (side * F(1e-3)) for k in range(3))
                for k in range(3):
                    tp[k] = np.where(alive, tp[k] * albedo[k], tp[k])
                o = tuple(np.where(alive, new_o[k], o[k]).astype(np.float32) for k in range(3))
                d = tuple(np.where(alive, new_d[k], d[k]).astype(np.float32) for k in range(3))
                depth += 1
            for k in range(3):                                                    # step 4
                accum[:, k] = accum[:, k] + L[k]
            accum[:, 3] = accum[:, 3] + F(1.0)
    return dict(accum=accum.reshape(height, width, 4), segments=segments)
