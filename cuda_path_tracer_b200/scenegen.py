"""Procedural scenes in the reference's own formats (scene.json + OBJ [+ MTL, textures, .hdr]).

The reference ships no assets (its scenes/ directory is git-ignored) and this box has no network,
so BASELINE.json's configs are rendered on deterministic procedural stand-ins with the same
triangle counts and material mix.  Every generator is a pure function of its arguments: the
reference build (oracle/_ref/refdump) and this library both load the files written here.
Every mesh carries `vn` records: a vertex without a normal becomes NaN at upload in the
reference (src/wavefront_internal.cu:139).
"""
import json
import os
import struct

import numpy as np


# ------------------------------------------------------------------------------------ primitives
def grid(origin, du, dv, nu, nv, normal, bump=0.0, seed=0):
    """(nu x nv) quads spanning origin + s*du + t*dv, displaced along `normal` by smooth noise."""
    origin, du, dv, normal = (np.asarray(a, np.float64) for a in (origin, du, dv, normal))
    s, t = np.meshgrid(np.linspace(0, 1, nu + 1), np.linspace(0, 1, nv + 1), indexing="ij")
    p = origin + s[..., None] * du + t[..., None] * dv
    n = np.broadcast_to(normal, p.shape).copy()
    if bump > 0:
        rng = np.random.RandomState(seed)
        ph = rng.uniform(0, 2 * np.pi, 6)
        fr = rng.uniform(3.0, 11.0, 6)
        h = bump * (np.sin(fr[0] * s * 2 * np.pi + ph[0]) * np.cos(fr[1] * t * 2 * np.pi + ph[1])
                    + 0.5 * np.sin(fr[2] * (s + t) * 2 * np.pi + ph[2]) + 0.25 * np.cos(fr[3] * (s - t) * 2 * np.pi + ph[3]))
        p = p + h[..., None] * normal
        dhs = np.gradient(h, axis=0) * nu
        dht = np.gradient(h, axis=1) * nv
        ts = du + dhs[..., None] * normal
        tt = dv + dht[..., None] * normal
        n = np.cross(ts, tt)
        n /= np.linalg.norm(n, axis=-1, keepdims=True)
        n *= np.sign((n * normal).sum(-1, keepdims=True))
    idx = np.arange((nu + 1) * (nv + 1)).reshape(nu + 1, nv + 1)
    a, b, c, d = idx[:-1, :-1], idx[1:, :-1], idx[1:, 1:], idx[:-1, 1:]
    flip = np.dot(np.cross(du, dv), normal) < 0
    t1 = np.stack([a, b, c], -1) if not flip else np.stack([a, c, b], -1)
    t2 = np.stack([a, c, d], -1) if not flip else np.stack([a, d, c], -1)
    tris = np.concatenate([t1.reshape(-1, 3), t2.reshape(-1, 3)])
    uv = np.stack([s, t], -1).reshape(-1, 2)
    return p.reshape(-1, 3), n.reshape(-1, 3), uv, tris


def sphere(center, radius, stacks, slices):
    """UV sphere, 2*slices*(stacks-1) triangles."""
    center = np.asarray(center, np.float64)
    th = np.linspace(0, np.pi, stacks + 1)[1:-1]
    ph = np.linspace(0, 2 * np.pi, slices, endpoint=False)
    T, Ph = np.meshgrid(th, ph, indexing="ij")
    n = np.stack([np.sin(T) * np.cos(Ph), np.cos(T), np.sin(T) * np.sin(Ph)], -1).reshape(-1, 3)
    n = np.concatenate([[[0, 1, 0]], n, [[0, -1, 0]]])
    p = center + radius * n
    uv = np.concatenate([[[0.5, 0.0]], np.stack([Ph / (2 * np.pi), T / np.pi], -1).reshape(-1, 2), [[0.5, 1.0]]])
    tris = []
    ring = lambda r: 1 + r * slices + np.arange(slices)
    r0 = ring(0)
    tris.append(np.stack([np.zeros(slices, int), np.roll(r0, -1), r0], -1))
    for r in range(stacks - 2):
        a, b = ring(r), ring(r + 1)
        a2, b2 = np.roll(a, -1), np.roll(b, -1)
        tris.append(np.stack([a, a2, b2], -1))
        tris.append(np.stack([a, b2, b], -1))
    rl = ring(stacks - 2)
    last = 1 + (stacks - 1) * slices
    tris.append(np.stack([rl, np.roll(rl, -1), np.full(slices, last)], -1))
    return p, n, uv, np.concatenate(tris)


def box(lo, hi):
    """Axis-aligned box, 12 triangles, flat normals, outward."""
    lo, hi = np.asarray(lo, np.float64), np.asarray(hi, np.float64)
    P, N, U, T = [], [], [], []
    for axis in range(3):
        for sign in (-1, 1):
            u, v = (axis + 1) % 3, (axis + 2) % 3
            c = np.zeros((4, 3))
            c[:, axis] = hi[axis] if sign > 0 else lo[axis]
            c[:, u] = [lo[u], hi[u], hi[u], lo[u]]
            c[:, v] = [lo[v], lo[v], hi[v], hi[v]]
            n = np.zeros(3)
            n[axis] = sign
            b = 4 * len(T) // 2
            P.append(c)
            N.append(np.tile(n, (4, 1)))
            U.append(np.array([[0, 0], [1, 0], [1, 1], [0, 1]], float))
            T.append(np.array([[b, b + 1, b + 2], [b, b + 2, b + 3]]) if sign > 0 else np.array([[b, b + 2, b + 1], [b, b + 3, b + 2]]))
    return np.concatenate(P), np.concatenate(N), np.concatenate(U), np.concatenate(T)


def ribbon(start, step, up, normal, count):
    """A strip of exactly `count` triangles (filler to reach an exact triangle budget)."""
    start, step, up, normal = (np.asarray(a, np.float64) for a in (start, step, up, normal))
    k = count + 2
    i = np.arange(k)
    p = start + (i // 2)[:, None] * step + (i % 2)[:, None] * up
    n = np.tile(normal / np.linalg.norm(normal), (k, 1))
    uv = np.stack([(i // 2) / max(1, k // 2), (i % 2).astype(float)], -1)
    t = np.stack([i[:-2], i[1:-1], i[2:]], -1)
    odd = (np.arange(count) % 2) == 1
    t[odd] = t[odd][:, [1, 0, 2]]
    if np.dot(np.cross(step, up), normal) < 0:
        t = t[:, [0, 2, 1]]
    return p, n, uv, t


def merge(parts):
    P, N, U, T, base = [], [], [], [], 0
    for p, n, u, t in parts:
        P.append(p); N.append(n); U.append(u); T.append(t + base)
        base += len(p)
    return np.concatenate(P), np.concatenate(N), np.concatenate(U), np.concatenate(T)


# ------------------------------------------------------------------------------------ writers
def write_obj(path, p, n, uv, tris, mtllib=None, usemtl=None, with_uv=False):
    """One OBJ: v / vn [/ vt] per vertex, `f a/a/a` or `f a//a`.  usemtl: None or per-triangle names."""
    lines = []
    if mtllib:
        lines.append(f"mtllib {mtllib}")
    lines += ["v %.6f %.6f %.6f" % tuple(r) for r in p]
    lines += ["vn %.6f %.6f %.6f" % tuple(r) for r in n]
    if with_uv:
        lines += ["vt %.6f %.6f" % tuple(r) for r in uv]
    cur = None
    t1 = tris + 1
    for k, (a, b, c) in enumerate(t1):
        if usemtl is not None and usemtl[k] != cur:
            cur = usemtl[k]
            lines.append(f"usemtl {cur}")
        if with_uv:
            lines.append(f"f {a}/{a}/{a} {b}/{b}/{b} {c}/{c}/{c}")
        else:
            lines.append(f"f {a}//{a} {b}//{b} {c}//{c}")
    with open(path, "w") as f:
        f.write("\n".join(lines) + "\n")


def write_bmp(path, rgb):
    """24-bit uncompressed BMP (bottom-up), readable by stb_image and by csrc/host/image_io.cpp."""
    h, w, _ = rgb.shape
    row = (3 * w + 3) & ~3
    data = np.zeros((h, row), np.uint8)
    data[:, :3 * w] = rgb[::-1, :, ::-1].reshape(h, 3 * w)
    with open(path, "wb") as f:
        f.write(b"BM" + struct.pack("<IHHI", 54 + row * h, 0, 0, 54))
        f.write(struct.pack("<IiiHHIIiiII", 40, w, h, 1, 24, 0, row * h, 2835, 2835, 0, 0))
        f.write(data.tobytes())


def write_hdr(path, rgb):
    """Radiance .hdr, flat (non-RLE) RGBE scanlines."""
    h, w, _ = rgb.shape
    m = rgb.max(-1)
    e = np.where(m > 1e-32, np.floor(np.log2(np.maximum(m, 1e-38))) + 1, 0)
    scale = np.where(m > 1e-32, 256.0 / np.exp2(e), 0.0)
    out = np.zeros((h, w, 4), np.uint8)
    out[..., :3] = np.clip(rgb * scale[..., None], 0, 255).astype(np.uint8)
    out[..., 3] = np.where(m > 1e-32, e + 128, 0).astype(np.uint8)
    with open(path, "wb") as f:
        f.write(b"#?RADIANCE\nFORMAT=32-bit_rle_rgbe\n\n" + f"-Y {h} +X {w}\n".encode())
        f.write(out.tobytes())


def write_scene(path, camera, materials, objects, env=None):
    doc = {"Camera": camera, "Materials": materials, "Objects": objects}
    if env:
        doc["Environment Texture"] = {"FILE": env}
    with open(path, "w") as f:
        json.dump(doc, f, indent=1)


def _obj_entry(file, material=None):
    e = {"TRANS": [0, 0, 0], "ROTAT": [0, 0, 0], "SCALE": [1, 1, 1], "FILE": file}
    if material:
        e["MATERIAL"] = material
    return e


def _mat(color, kind="DIFFUSE", metallic=0.0, roughness=0.5, emittance=0.0, ior=1.0):
    return {"basecolor": list(color), "metallic": metallic, "roughness": roughness, "emittance": emittance, "ior": ior, "Type": kind}


# ------------------------------------------------------------------------------------ config 1
def cornell(out_dir, res=(800, 800), depth=8):
    """Cornell-style box in the reference's schema: Lambertian walls + one emissive quad + two boxes."""
    os.makedirs(out_dir, exist_ok=True)
    mats = {"white": _mat([0.73, 0.73, 0.73]), "red": _mat([0.65, 0.05, 0.05]), "green": _mat([0.12, 0.45, 0.15]),
            "light": _mat([1.0, 0.9, 0.7], emittance=12.0)}
    objs = []

    def add(name, part, material):
        write_obj(os.path.join(out_dir, name + ".obj"), *part)
        objs.append(_obj_entry(os.path.join(out_dir, name + ".obj"), material))

    add("floor", grid([-1, 0, -1], [2, 0, 0], [0, 0, 2], 2, 2, [0, 1, 0]), "white")
    add("ceiling", grid([-1, 2, -1], [2, 0, 0], [0, 0, 2], 2, 2, [0, -1, 0]), "white")
    add("back", grid([-1, 0, -1], [2, 0, 0], [0, 2, 0], 2, 2, [0, 0, 1]), "white")
    add("left", grid([-1, 0, -1], [0, 0, 2], [0, 2, 0], 2, 2, [1, 0, 0]), "red")
    add("right", grid([1, 0, -1], [0, 0, 2], [0, 2, 0], 2, 2, [-1, 0, 0]), "green")
    add("lamp", grid([-0.3, 1.995, -0.3], [0.6, 0, 0], [0, 0, 0.6], 1, 1, [0, -1, 0]), "light")
    add("tall", box([-0.65, 0.0, -0.6], [-0.1, 1.2, -0.05]), "white")
    add("short", box([0.1, 0.0, 0.0], [0.65, 0.6, 0.55]), "white")
    cam = {"RES": list(res), "FOVY": 39.0, "ITERATIONS": 16, "DEPTH": depth, "FILE": "cornell",
           "EYE": [0.0, 1.0, 3.9], "LOOKAT": [0.0, 1.0, 0.0], "UP": [0.0, 1.0, 0.0]}
    path = os.path.join(out_dir, "scene.json")
    write_scene(path, cam, mats, objs)
    return path


# ------------------------------------------------------------------------------------ config 2
FIREPLACE_MIX = {"MicrofacetPBR": 24421, "DIFFUSE": 107572, "SPECULAR_REFLECTION": 10, "SPECULAR_REFRACTION": 11170}


def fireplace(out_dir, res=(1920, 1080), depth=8, scale=1.0):
    """143,173-triangle room with the material mix the reference README quotes for Fireplace Room
    (README.md:34): PBR 24,421 / diffuse 107,572 / mirror 10 / dielectric 11,170.  `scale` < 1
    shrinks every tessellation (for quick tests); the mix is exact only at scale == 1."""
    os.makedirs(out_dir, exist_ok=True)
    rng = np.random.RandomState(20240607)
    q = lambda v: max(2, int(round(v * scale)))
    mats = {
        "wall": _mat([0.72, 0.68, 0.62]), "floor": _mat([0.45, 0.33, 0.25]), "brick": _mat([0.55, 0.22, 0.16]),
        "ball": _mat([0.25, 0.35, 0.65]), "trim": _mat([0.8, 0.8, 0.75]),
        "lamp": _mat([1.0, 0.85, 0.6], emittance=18.0),
        "gold": _mat([1.0, 0.77, 0.34], "MicrofacetPBR", metallic=1.0, roughness=0.25),
        "plastic": _mat([0.8, 0.1, 0.1], "MicrofacetPBR", metallic=0.0, roughness=0.35),
        "steel": _mat([0.6, 0.62, 0.65], "MicrofacetPBR", metallic=0.9, roughness=0.5),
        "ceramic": _mat([0.9, 0.9, 0.85], "MicrofacetPBR", metallic=0.0, roughness=0.15),
        "glass": _mat([0.95, 0.98, 1.0], "SPECULAR_REFRACTION", ior=1.5),
        "mirror": _mat([0.95, 0.95, 0.95], "SPECULAR_REFLECTION", metallic=1.0, roughness=0.0),
    }
    objs, counts = [], {k: 0 for k in FIREPLACE_MIX}

    def add(name, part, material):
        write_obj(os.path.join(out_dir, name + ".obj"), *part)
        objs.append(_obj_entry(os.path.join(out_dir, name + ".obj"), material))
        counts[mats[material]["Type"]] += len(part[3])

    X, Y, Z = 5.0, 5.0, 5.0
    g = q(100)
    add("floor", grid([-X, 0, -Z], [2 * X, 0, 0], [0, 0, 2 * Z], g, g, [0, 1, 0], 0.03, 1), "floor")
    add("ceiling", grid([-X, Y, -Z], [2 * X, 0, 0], [0, 0, 2 * Z], g, g, [0, -1, 0], 0.03, 2), "wall")
    h = q(50)
    add("wall_back", grid([-X, 0, -Z], [2 * X, 0, 0], [0, Y, 0], g, h, [0, 0, 1], 0.04, 3), "wall")
    add("wall_front", grid([-X, 0, Z], [2 * X, 0, 0], [0, Y, 0], g, h, [0, 0, -1], 0.04, 4), "wall")
    add("wall_left", grid([-X, 0, -Z], [0, 0, 2 * Z], [0, Y, 0], g, h, [1, 0, 0], 0.04, 5), "wall")
    add("wall_right", grid([X, 0, -Z], [0, 0, 2 * Z], [0, Y, 0], g, h, [-1, 0, 0], 0.04, 6), "wall")
    bricks = []
    for _ in range(q(200)):   # a fireplace-ish pile of bricks along the back wall
        c = np.array([rng.uniform(-2.0, 2.0), rng.uniform(0.05, 2.2), rng.uniform(-4.7, -3.9)])
        s = np.array([rng.uniform(0.08, 0.2), rng.uniform(0.04, 0.09), rng.uniform(0.06, 0.12)])
        bricks.append(box(c - s, c + s))
    add("bricks", merge(bricks), "brick")
    balls = [sphere([rng.uniform(-3.5, 3.5), 0.45, rng.uniform(-2.5, 2.5)], 0.42, q(40), q(60)) for _ in range(5)]
    add("balls", merge(balls), "ball")
    for k, m in enumerate(["gold", "plastic", "steel", "ceramic"]):
        add(f"pbr_{m}", sphere([-3.0 + 2.0 * k, 1.0 + 0.25 * (k % 2), -1.0 - 0.6 * (k % 2)], 0.7, q(48), q(64)), m)
    glass = [sphere([-1.2, 0.62, 1.4], 0.6, q(59), q(48)), sphere([1.4, 0.52, 1.9], 0.5, q(59), q(48))]
    add("glass", merge(glass), "glass")
    lamps = [grid([cx - 0.5, Y - 0.06, cz - 0.5], [1.0, 0, 0], [0, 0, 1.0], 1, 1, [0, -1, 0]) for cx in (-2.2, 2.2) for cz in (-2.2, 2.2)]
    add("lamps", merge(lamps), "lamp")
    add("mirror", ribbon([X - 0.08, 1.0, -1.5], [0, 0, 0.6], [0, 2.0, 0], [-1, 0, 0], 10), "mirror")
    if scale == 1.0:   # fillers: bring each material class to the README's exact count
        add("trim_pbr", ribbon([-X + 0.05, 0.4, -4.0], [0, 0, 0.04], [0, 0.12, 0], [1, 0, 0], FIREPLACE_MIX["MicrofacetPBR"] - counts["MicrofacetPBR"]), "steel")
        add("trim_glass", ribbon([-2.0, 2.6, -Z + 0.06], [0.1, 0, 0], [0, 0.3, 0], [0, 0, 1], FIREPLACE_MIX["SPECULAR_REFRACTION"] - counts["SPECULAR_REFRACTION"]), "glass")
        add("trim_diffuse", ribbon([-X + 0.1, 0.08, Z - 0.06], [0.011, 0, 0], [0, 0.1, 0], [0, 0, -1], FIREPLACE_MIX["DIFFUSE"] - counts["DIFFUSE"]), "trim")
        assert counts == FIREPLACE_MIX, counts
    cam = {"RES": list(res), "FOVY": 55.0, "ITERATIONS": 16, "DEPTH": depth, "FILE": "fireplace",
           "EYE": [0.3, 2.2, 4.6], "LOOKAT": [0.0, 1.4, -2.0], "UP": [0.0, 1.0, 0.0]}
    path = os.path.join(out_dir, "scene.json")
    write_scene(path, cam, mats, objs)
    return path, sum(counts.values())


# ------------------------------------------------------------------------------------ config 4
def heightfield_arrays(num_triangles, seed=7, extent=40.0):
    """Displaced height field + scattered micro-triangles as plain arrays (too large for OBJ text):
    pos (V,3), nor (V,3), indices (N,3), material_ids (N,), all for Scene.from_arrays."""
    n_field = int(num_triangles * 0.9)
    g = int(np.sqrt(n_field / 2))
    p, n, uv, t = grid([-extent, 0, -extent], [2 * extent, 0, 0], [0, 0, 2 * extent], g, g, [0, 1, 0], 1.5, seed)
    rest = num_triangles - len(t)
    rng = np.random.RandomState(seed)
    c = np.stack([rng.uniform(-extent, extent, rest), rng.uniform(2.0, 12.0, rest), rng.uniform(-extent, extent, rest)], -1)
    d1 = rng.normal(0, 0.08, (rest, 3))
    d2 = rng.normal(0, 0.08, (rest, 3))
    mp = np.stack([c, c + d1, c + d2], 1).reshape(-1, 3)
    mn = np.cross(d1, d2)
    mn /= np.maximum(np.linalg.norm(mn, axis=-1, keepdims=True), 1e-20)
    mn = np.repeat(mn, 3, axis=0)
    mt = np.arange(3 * rest).reshape(-1, 3) + len(p)
    pos = np.concatenate([p, mp]).astype(np.float32)
    nor = np.concatenate([n, mn]).astype(np.float32)
    idx = np.concatenate([t, mt]).astype(np.int32)
    mat = np.zeros(len(idx), np.int32)
    mat[len(t):] = 1 + (np.arange(rest) % 3)
    # a few large emissive triangles high above the field
    lp = np.array([[-6, 30, -6], [6, 30, -6], [6, 30, 6], [-6, 30, 6]], np.float32)
    ln = np.tile(np.array([[0, -1, 0]], np.float32), (4, 1))
    base = len(pos)
    pos = np.concatenate([pos, lp]); nor = np.concatenate([nor, ln])
    idx[-2] = [base, base + 1, base + 2]; idx[-1] = [base, base + 2, base + 3]
    mat[-2:] = 4
    return pos, nor, idx, mat


def _fixed_floats(a, int_digits=4, frac_digits=6):
    """(n, k) floats -> (n, k * width) ASCII bytes ' -dddd.dddddd' per value (fixed width, leading zeros), vectorised."""
    a = np.asarray(a, np.float64)
    q = np.rint(np.abs(a) * 10 ** frac_digits).astype(np.int64)
    width = 2 + int_digits + 1 + frac_digits
    out = np.empty(a.shape + (width,), np.uint8)
    out[..., 0] = ord(" ")
    out[..., 1] = np.where((a < 0) & (q > 0), ord("-"), ord("+"))
    col = width - 1
    for k in range(frac_digits):
        out[..., col] = ord("0") + q % 10; q //= 10; col -= 1
    out[..., col] = ord("."); col -= 1
    for k in range(int_digits):
        out[..., col] = ord("0") + q % 10; q //= 10; col -= 1
    return out.reshape(a.shape[0], -1)


def write_obj_fast(path, p, n, tris):
    """The same OBJ dialect as write_obj (v / vn per vertex, `f a//a b//b c//c`) in fixed-width fields built with numpy, for
    meshes of millions of triangles (10 M triangles: ~0.9 GB of text in a few seconds instead of minutes)."""
    nv, nt = len(p), len(tris)
    with open(path, "wb") as f:
        for tag, arr in ((b"v", p), (b"vn", n)):
            body = _fixed_floats(arr, 4 if tag == b"v" else 1)
            line = np.empty((nv, len(tag) + body.shape[1] + 1), np.uint8)
            line[:, :len(tag)] = np.frombuffer(tag, np.uint8)
            line[:, len(tag):-1] = body
            line[:, -1] = ord("\n")
            f.write(line.tobytes())
        t1 = np.asarray(tris, np.int64) + 1
        digits = 8
        idx = np.empty((nt, 3, digits), np.uint8)
        q = t1.copy()
        for k in range(digits - 1, -1, -1):
            idx[..., k] = ord("0") + q % 10; q //= 10
        # "f A//A B//B C//C\n"
        line = np.empty((nt, 1 + 3 * (1 + digits + 2 + digits) + 1), np.uint8)
        line[:, 0] = ord("f")
        c = 1
        for v in range(3):
            line[:, c] = ord(" "); c += 1
            line[:, c:c + digits] = idx[:, v]; c += digits
            line[:, c] = ord("/"); line[:, c + 1] = ord("/"); c += 2
            line[:, c:c + digits] = idx[:, v]; c += digits
        line[:, c] = ord("\n")
        f.write(line.tobytes())


def stress_scene(out_dir, num_triangles=10_000_000, res=(1920, 1080), depth=8):
    """BASELINE config[3] in scene-file form (what the reference's own loader can read): the height field + micro-triangles of
    heightfield_arrays() written as one OBJ per material, the five stress materials and the stress camera."""
    os.makedirs(out_dir, exist_ok=True)
    pos, nor, idx, mat = heightfield_arrays(num_triangles)
    names = ["a_ground", "b_gold", "c_glass", "d_blue", "e_lamp"]
    spec = {"a_ground": _mat([0.55, 0.5, 0.42]), "b_gold": _mat([0.9, 0.6, 0.2], "MicrofacetPBR", metallic=1.0, roughness=0.3),
            "c_glass": _mat([0.95, 0.97, 1.0], "SPECULAR_REFRACTION", ior=1.5), "d_blue": _mat([0.2, 0.5, 0.8]),
            "e_lamp": _mat([1.0, 0.95, 0.85], emittance=40.0)}
    objs = []
    for m, name in enumerate(names):
        t = idx[mat == m]
        used, inv = np.unique(t.reshape(-1), return_inverse=True)
        path = os.path.join(out_dir, name + ".obj")
        write_obj_fast(path, pos[used], nor[used], inv.reshape(-1, 3))
        objs.append(_obj_entry(path, name))
    c = stress_camera(res)
    cam = {"RES": list(res), "FOVY": c["fovy"], "ITERATIONS": 4, "DEPTH": depth, "FILE": "stress", "EYE": c["eye"], "LOOKAT": c["look_at"], "UP": c["up"]}
    path = os.path.join(out_dir, "scene.json")
    write_scene(path, cam, spec, objs)
    return path, len(idx)


def stress_materials():
    from .capi import MATERIAL_DTYPE
    m = np.zeros(5, MATERIAL_DTYPE)
    spec = [([0.55, 0.5, 0.42], 1, 0, 0.5, 0, 1), ([0.9, 0.6, 0.2], 0, 1.0, 0.3, 0, 1), ([0.95, 0.97, 1.0], 3, 0, 0, 0, 1.5),
            ([0.2, 0.5, 0.8], 1, 0, 0.5, 0, 1), ([1.0, 0.95, 0.85], 1, 0, 0.5, 40.0, 1)]
    for i, (c, ty, me, ro, em, ior) in enumerate(spec):
        m[i]["basecolor"] = c; m[i]["type"] = ty; m[i]["metallic"] = me; m[i]["roughness"] = ro
        m[i]["emittance"] = em; m[i]["ior"] = ior
        m[i]["diffuse_tex_id"] = m[i]["normal_tex_id"] = m[i]["metallic_roughness_tex_id"] = -1
    return m


def stress_camera(res=(1920, 1080)):
    return dict(resolution=res, fovy=50.0, eye=[0.0, 14.0, 38.0], look_at=[0.0, 2.0, 0.0], up=[0.0, 1.0, 0.0])


# ------------------------------------------------------------------------------------ deep LBVH: the 32-entry stack abort
def deepchain(out_dir, res=(160, 120), depth=4, fovy=70.0):
    """A scene whose LBVH is a CHAIN deeper than the 32 entries the reference's closest-hit walk stacks before it gives up
    (src/kernels/ray_cast.cu:35,106), arranged so that camera rays really run into that limit.

    Keys are (30-bit Morton code << 32) | triangle index (src/bvh.cu:95) and Karras' hierarchy splits at the highest
    differing bit, so a chain needs keys that differ from one anchor key in exactly one bit each:
      * three slivers along the axes pin the scene box to [0, 1]^3 (codes are then exact binary fractions);
      * Q = (1/2, 1/2, 1/2) has code 111 000 ... 0; "marker" k (k = 0..26) is one triangle whose AABB centre is Q plus
        2^-(2 + k/3) on axis k % 3: its code is Q's with bit 26 - k set -- 27 chain levels from the code bits;
      * a cluster of triangles centred in Q's own cell takes the triangle indices 0, 1, 2, 4, ..., 256: identical codes,
        so the index bits continue the chain for 9 more levels.
    The camera sits at Q looking along (1, 1, 1): it is inside every "rest of the chain" box (entry distance 0) and just
    outside most marker boxes, which lie ahead and are all crossed by rays near the diagonal -- so the walk descends
    near-first into the chain and stacks one marker per level.  Markers 3..5 contain the eye (a tie: the
    right child, the marker, is tested first), so some rays give up holding a hit, others holding none."""
    os.makedirs(out_dir, exist_ok=True)
    Q = 0.5
    tris = {}                                   # triangle index -> (3, 3) vertices
    kind = {}                                   # triangle index -> material name

    def corner_tri(lo, hi):
        lo, hi = np.asarray(lo, np.float64), np.asarray(hi, np.float64)
        return np.array([[lo[0], hi[1], hi[2]], [hi[0], lo[1], hi[2]], [hi[0], hi[1], lo[2]]])   # the far diagonal of the box

    def axis_box(axis, s, eye_inside):
        lo = np.full(3, Q - 0.9 * s); hi = np.full(3, Q + 0.9 * s)
        lo[axis] = Q + (-0.1 if eye_inside else 0.1) * s
        hi[axis] = 2.0 * (Q + s) - lo[axis]      # AABB centre on this axis = Q + s exactly
        return lo, hi

    cluster_ids = [0, 1, 2, 4, 8, 16, 32, 64, 128, 256]
    cell = 1.0 / 1024.0
    for j, tid in enumerate(cluster_ids):        # centres inside Q's cell: same 30-bit code, order decided by the index
        s = cell * 0.45 * (j + 1) / len(cluster_ids)
        lo, hi = axis_box(j % 3, s, False)
        tris[tid] = corner_tri(lo, hi); kind[tid] = "white"
    free = [i for i in range(320) if i not in tris]
    it = iter(free)
    for k in range(27):
        s = 2.0 ** -(2 + k // 3)
        lo, hi = axis_box(k % 3, s, 3 <= k < 6)
        tid = next(it)
        tris[tid] = corner_tri(lo, hi); kind[tid] = "lamp" if k == 10 else ("red" if k % 2 else "white")   # a mid-chain marker lights the larger ones behind it
    eps = 1.0 / 4096.0
    for a in range(3):                           # slivers along the axes: scene box = [0, 1]^3, centres outside Q's octant
        v = np.zeros((3, 3)); v[1, a] = 1.0; v[2, a] = 0.5; v[2, (a + 1) % 3] = eps; v[1, (a + 2) % 3] = eps
        tid = next(it)
        tris[tid] = v; kind[tid] = "white"
    rng = np.random.RandomState(11)
    for tid in it:                               # filler in the x < 1/2 half: small triangles the camera never sees
        c = rng.uniform([0.05, 0.05, 0.05], [0.4, 0.9, 0.9])
        tris[tid] = c + rng.uniform(-0.02, 0.02, (3, 3)); kind[tid] = "green"
    n = len(tris)
    P = np.concatenate([tris[i] for i in range(n)]).astype(np.float64)
    Nn = np.repeat(np.array([np.cross(tris[i][1] - tris[i][0], tris[i][2] - tris[i][0]) for i in range(n)]), 3, axis=0)
    Nn /= np.maximum(np.linalg.norm(Nn, axis=1, keepdims=True), 1e-30)
    T = np.arange(3 * n).reshape(-1, 3)
    # %.6f would move the corners: write the vertices with enough digits to round-trip float32
    lines = ["v %.9g %.9g %.9g" % tuple(np.float32(r)) for r in P] + ["vn %.6f %.6f %.6f" % tuple(r) for r in Nn]
    cur = None
    for i, (a, b, c) in enumerate(T + 1):
        if kind[i] != cur:
            cur = kind[i]
            lines.append(f"usemtl {cur}")
        lines.append(f"f {a}//{a} {b}//{b} {c}//{c}")
    mtl = {"white": [0.73, 0.73, 0.73], "red": [0.65, 0.05, 0.05], "green": [0.12, 0.45, 0.15], "lamp": [1.0, 0.9, 0.7]}
    with open(os.path.join(out_dir, "chain.mtl"), "w") as f:
        for name, kd in mtl.items():
            f.write(f"newmtl {name}\nKd {kd[0]} {kd[1]} {kd[2]}\n" + ("Ke 14 12 9\n" if name == "lamp" else "") + "\n")
    with open(os.path.join(out_dir, "chain.obj"), "w") as f:
        f.write("mtllib chain.mtl\n" + "\n".join(lines) + "\n")
    cam = {"RES": list(res), "FOVY": fovy, "ITERATIONS": 4, "DEPTH": depth, "FILE": "deepchain",
           "EYE": [Q, Q, Q], "LOOKAT": [1.0, 1.0, 1.0], "UP": [0.0, 1.0, 0.0]}
    path = os.path.join(out_dir, "scene.json")
    write_scene(path, cam, {"unused": _mat([0.5, 0.5, 0.5])}, [_obj_entry(os.path.join(out_dir, "chain.obj"))])
    return path


# ------------------------------------------------------------------------------------ config 3
def cylinder(base, radius, height, slices, stacks):
    """Open cylinder along +y, 2*slices*stacks triangles, seam duplicated so that u runs 0..1."""
    base = np.asarray(base, np.float64)
    a, t = np.meshgrid(np.linspace(0, 2 * np.pi, slices + 1), np.linspace(0, 1, stacks + 1), indexing="ij")
    n = np.stack([np.cos(a), np.zeros_like(a), np.sin(a)], -1)
    p = base + radius * n + np.stack([np.zeros_like(t), t * height, np.zeros_like(t)], -1)
    idx = np.arange((slices + 1) * (stacks + 1)).reshape(slices + 1, stacks + 1)
    q0, q1, q2, q3 = idx[:-1, :-1], idx[1:, :-1], idx[1:, 1:], idx[:-1, 1:]
    tris = np.concatenate([np.stack([q0, q2, q1], -1).reshape(-1, 3), np.stack([q0, q3, q2], -1).reshape(-1, 3)])
    uv = np.stack([a / (2 * np.pi) * 3.0, t * 4.0], -1).reshape(-1, 2)
    return p.reshape(-1, 3), n.reshape(-1, 3), uv, tris


def vault(x0, x1, z, radius, y0, nu, nv):
    """Half-cylinder arch spanning x0..x1 at depth z..z+1.2, facing down (normals toward the axis)."""
    a, s = np.meshgrid(np.linspace(0, np.pi, nu + 1), np.linspace(0, 1, nv + 1), indexing="ij")
    cx = 0.5 * (x0 + x1)
    n = -np.stack([np.cos(a), np.sin(a), np.zeros_like(a)], -1)
    p = np.stack([cx + radius * np.cos(a), y0 + radius * np.sin(a), z + 1.2 * s], -1)
    idx = np.arange((nu + 1) * (nv + 1)).reshape(nu + 1, nv + 1)
    q0, q1, q2, q3 = idx[:-1, :-1], idx[1:, :-1], idx[1:, 1:], idx[:-1, 1:]
    tris = np.concatenate([np.stack([q0, q1, q2], -1).reshape(-1, 3), np.stack([q0, q2, q3], -1).reshape(-1, 3)])
    uv = np.stack([a / np.pi * 2.0, s], -1).reshape(-1, 2)
    return p.reshape(-1, 3), n.reshape(-1, 3), uv, tris


def _noise(size, seed, octaves=4):
    """Tileable value noise in [0,1] (sum of random-phase sinusoid products; deterministic)."""
    rng = np.random.RandomState(seed)
    y, x = np.meshgrid(np.linspace(0, 2 * np.pi, size, endpoint=False), np.linspace(0, 2 * np.pi, size, endpoint=False), indexing="ij")
    h = np.zeros((size, size))
    for o in range(octaves):
        for _ in range(3):
            fx, fy = rng.randint(1, 4 << o, 2)
            h += np.sin(fx * x + rng.uniform(0, 6.28)) * np.cos(fy * y + rng.uniform(0, 6.28)) / (1 << o)
    h -= h.min()
    return h / h.max()


def _normal_map(height, strength):
    gx = (np.roll(height, -1, 1) - np.roll(height, 1, 1)) * strength
    gy = (np.roll(height, -1, 0) - np.roll(height, 1, 0)) * strength
    n = np.stack([-gx, -gy, np.ones_like(gx)], -1)
    n /= np.linalg.norm(n, axis=-1, keepdims=True)
    return np.clip((n * 0.5 + 0.5) * 255, 0, 255).astype(np.uint8)


def _texture_set(out_dir, name, size, seed, tint, pattern):
    """diffuse / normal / metallic-roughness BMPs for one surface; returns the three file names."""
    h = _noise(size, seed)
    yy, xx = np.mgrid[0:size, 0:size]
    if pattern == "brick":
        row = (yy * 16 // size)
        mortar = ((yy * 16 % size) < size // 40) | ((((xx + (row % 2) * size // 16) * 8) % size) < size // 40)
        h = np.where(mortar, 0.15, 0.55 + 0.45 * h)
    elif pattern == "tile":
        mortar = ((yy * 8 % size) < size // 64) | ((xx * 8 % size) < size // 64)
        h = np.where(mortar, 0.1, 0.6 + 0.4 * h * (((yy * 8 // size) + (xx * 8 // size)) % 2 * 0.5 + 0.5))
    elif pattern == "cloth":
        h = 0.5 + 0.25 * np.sin(xx * 2 * np.pi * 64 / size) * np.sin(yy * 2 * np.pi * 64 / size) + 0.25 * h
    col = np.clip(np.asarray(tint)[None, None, :] * (0.35 + 0.65 * h[..., None]) * 255, 0, 255).astype(np.uint8)
    files = [f"{name}_diffuse.bmp", f"{name}_normal.bmp", f"{name}_rough.bmp"]
    write_bmp(os.path.join(out_dir, files[0]), col)
    write_bmp(os.path.join(out_dir, files[1]), _normal_map(h, 6.0))
    mr = np.zeros((size, size, 3), np.uint8)
    mr[..., 1] = np.clip((0.35 + 0.6 * _noise(size, seed + 100, 3)) * 255, 0, 255)   # roughness factor (green)
    mr[..., 2] = np.clip((pattern == "tile") * 0.5 * 255 * _noise(size, seed + 200, 2), 0, 255)   # metallic factor (blue)
    write_bmp(os.path.join(out_dir, files[2]), mr)
    return files


def sky_hdr(path, width=2048, height=1024, sun_dir=(0.35, 0.8, 0.45), sun_radiance=4000.0):
    """Equirect RGB32F sky: horizon-to-zenith gradient + a small bright sun (exercises the alias table and the
    pdf-map MIS of src/kernels/logic.cu:76-103)."""
    v, u = np.meshgrid((np.arange(height) + 0.5) / height, (np.arange(width) + 0.5) / width, indexing="ij")
    theta, phi = v * np.pi, u * 2 * np.pi
    d = np.stack([np.sin(theta) * np.cos(phi), np.cos(theta), np.sin(theta) * np.sin(phi)], -1)
    up = np.clip(d[..., 1], 0, 1)
    sky = np.stack([0.35 + 0.25 * (1 - up), 0.5 + 0.2 * (1 - up), 0.9 * np.ones_like(up)], -1) * (0.6 + 0.8 * up[..., None])
    sky = np.where(d[..., 1:2] < 0, np.array([0.12, 0.1, 0.08]), sky)
    s = np.asarray(sun_dir, np.float64)
    s /= np.linalg.norm(s)
    cosang = (d * s).sum(-1)
    sky = sky + (cosang > np.cos(np.radians(1.5)))[..., None] * np.array([1.0, 0.93, 0.8]) * sun_radiance
    write_hdr(path, sky.astype(np.float32))


SPONZA_TRIANGLES = 262279


def sponza(out_dir, res=(1920, 1080), depth=8, scale=1.0, tex_size=1024):
    """262,279-triangle atrium (README.md:115 quotes Sponza at that size): textured floor / walls / columns /
    arches / drapes through OBJ + MTL (map_Kd, map_bump, map_Pr => bindless textures), emissive lanterns (Ke),
    open roof under an HDR sky with a sun.  scale < 1 shrinks the tessellation; the count is exact at scale 1."""
    os.makedirs(out_dir, exist_ok=True)
    q = lambda v: max(2, int(round(v * scale)))
    sets = {"floor": ("tile", [0.8, 0.75, 0.7], 11), "wall": ("brick", [0.75, 0.45, 0.35], 12), "column": ("noise", [0.85, 0.82, 0.75], 13),
            "arch": ("brick", [0.7, 0.68, 0.6], 14), "drape_red": ("cloth", [0.75, 0.12, 0.1], 15), "drape_blue": ("cloth", [0.15, 0.25, 0.7], 16)}
    mtl, tex_count = [], 0
    for name, (pattern, tint, seed) in sets.items():
        f = _texture_set(out_dir, name, tex_size, seed, tint, pattern)
        use = 3 if name in ("floor", "wall", "column") else 2        # 3+3+3+2+2+2 = 15 textures (+ env, + pdf map)
        tex_count += use
        mtl += [f"newmtl {name}", "Kd 0.9 0.9 0.9", "Ks 0.04 0.04 0.04", "Ns 400", "Ni 1.0", "d 1.0", "illum 2", f"map_Kd {f[0]}", f"map_bump {f[1]}"]
        if use == 3:
            mtl.append(f"map_Pr {f[2]}")
        mtl.append("")
    mtl += ["newmtl plaster", "Kd 0.8 0.78 0.7", "Ks 0 0 0", "Ns 10", "Ni 1.0", "d 1.0", "illum 2", "",
            "newmtl lantern", "Kd 1 1 1", "Ke 30 24 15", "Ks 0 0 0", "Ns 10", "Ni 1.0", "d 1.0", "illum 2", ""]
    with open(os.path.join(out_dir, "atrium.mtl"), "w") as fh:
        fh.write("\n".join(mtl))
    objs, total = [], [0]

    def add(name, part, material):
        names = material if isinstance(material, list) else [material] * len(part[3])
        write_obj(os.path.join(out_dir, name + ".obj"), *part, mtllib="atrium.mtl", usemtl=names, with_uv=True)
        objs.append(_obj_entry(os.path.join(out_dir, name + ".obj")))
        total[0] += len(part[3])

    X, Y, Z = 6.0, 9.0, 16.0
    g = q(160)
    add("floor", grid([-X, 0, -Z], [2 * X, 0, 0], [0, 0, 2 * Z], g, g, [0, 1, 0], 0.02, 21), "floor")
    add("wall_left", grid([-X, 0, -Z], [0, 0, 2 * Z], [0, Y, 0], g, q(60), [1, 0, 0], 0.05, 22), "wall")
    add("wall_right", grid([X, 0, -Z], [0, 0, 2 * Z], [0, Y, 0], g, q(60), [-1, 0, 0], 0.05, 23), "wall")
    add("wall_back", grid([-X, 0, -Z], [2 * X, 0, 0], [0, Y, 0], q(80), q(60), [0, 0, 1], 0.05, 24), "wall")
    add("wall_front", grid([-X, 0, Z], [2 * X, 0, 0], [0, Y, 0], q(80), q(60), [0, 0, -1], 0.05, 25), "plaster")
    cols, arches = [], []
    for k in range(10):
        z = -Z + 1.6 + k * (2 * Z - 3.2) / 9
        for x in (-3.4, 3.4):
            cols.append(cylinder([x, 0, z], 0.45, 5.0, q(48), q(40)))
        arches.append(vault(-3.4, 3.4, z - 0.6, 3.4, 5.0, q(48), q(24)))
    add("columns", merge(cols), "column")
    add("arches", merge(arches), "arch")
    drapes, names = [], []
    for k in range(6):
        z = -Z + 3.0 + k * 5.0
        x = -X + 0.35 if k % 2 == 0 else X - 0.35
        nrm = [1, 0, 0] if k % 2 == 0 else [-1, 0, 0]
        d = grid([x, 2.0, z], [0, 0, 2.4], [0, 5.0, 0], q(60), q(50), nrm, 0.12, 30 + k)
        drapes.append(d)
        names += ["drape_red" if k % 3 else "drape_blue"] * len(d[3])
    add("drapes", merge(drapes), names)           # one OBJ, two MTL materials: exercises the usemtl switch
    lanterns = [box([x - 0.15, 4.2, z - 0.15], [x + 0.15, 4.5, z + 0.15]) for x in (-2.0, 2.0) for z in (-8.0, 8.0)]
    add("lanterns", merge(lanterns), "lantern")
    if scale == 1.0:
        add("trim", ribbon([-X + 0.1, 0.05, -Z + 0.05], [0.0013, 0, 0], [0, 0.15, 0], [0, 0, 1], SPONZA_TRIANGLES - total[0]), "plaster")
        assert total[0] == SPONZA_TRIANGLES, total[0]
    sky_hdr(os.path.join(out_dir, "sky.hdr"), 2048 if scale == 1.0 else 512, 1024 if scale == 1.0 else 256)
    cam = {"RES": list(res), "FOVY": 60.0, "ITERATIONS": 16, "DEPTH": depth, "FILE": "atrium",
           "EYE": [0.4, 2.4, 13.5], "LOOKAT": [-0.3, 3.2, -4.0], "UP": [0.0, 1.0, 0.0]}
    path = os.path.join(out_dir, "scene.json")
    write_scene(path, cam, {}, objs, env=os.path.join(out_dir, "sky.hdr"))
    return path, total[0]
