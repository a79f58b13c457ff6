"""Scene descriptions for PtScene_NewFromDesc and the reference's procedural scenes.

cornell_box() restates CmdCornellBox / GenBoxMesh / GenSphereMesh / GenMaterial
(/root/reference/src/rendering/render_system.c:745-1306) in float32 numpy. The
other generators build the synthetic benchmark inputs BASELINE.json names.
"""
from dataclasses import dataclass, field
from typing import List, Optional

import numpy as np

f32 = np.float32

MATFLAG_EMISSIVE = 1 << 0
MATFLAG_SKY = 1 << 1
MATFLAG_REFRACTIVE = 1 << 5

VK_FORMAT_R8G8B8A8_SRGB = 43
VK_FORMAT_R16G16_SNORM = 78
TEX_RGBA8 = 0
TEX_RG16 = 1


@dataclass
class Texture:
    width: int
    height: int
    texels: np.ndarray          # uint32 [h*w] (RGBA8 packed little endian, or 2 x int16)
    format: int = TEX_RGBA8

    @property
    def vk_format(self):
        return VK_FORMAT_R8G8B8A8_SRGB if self.format == TEX_RGBA8 else VK_FORMAT_R16G16_SNORM


@dataclass
class Entity:
    positions: np.ndarray       # float32 [n,4] object space, unindexed triangles
    normals: np.ndarray         # float32 [n,4]
    uvs: np.ndarray             # float32 [n,4]
    matrix: np.ndarray          # float32 [16] column major local-to-world
    albedo: int = -1            # texture indices
    rome: int = -1
    normal: int = -1
    flags: int = 0
    mean_free_path: tuple = (0.0, 0.0, 0.0, 0.0)
    ior: float = 1.0
    bumpiness: float = 0.0


@dataclass
class SceneDesc:
    entities: List[Entity] = field(default_factory=list)
    textures: List[Texture] = field(default_factory=list)
    sky: Optional[np.ndarray] = None        # float32 [6, size, size, 3]
    brdf_lut: Optional[np.ndarray] = None   # float32 [h, w, 2]
    pt_dist_meters: float = 1.5
    seed: int = 1
    path_pool_size: int = 0
    bake_mode: int = 0              # 0 wavefront, 1 sequential (texel-exact vs the seeded reference)

    @property
    def tri_count(self):
        return sum(len(e.positions) for e in self.entities) // 3


# ----------------------------------------------------------------- float32 math

def _v(*a):
    return np.array(a, dtype=f32)


def _normalize3(v):
    v = v.astype(f32)
    l = f32(np.sqrt(f32(v[0] * v[0] + v[1] * v[1] + v[2] * v[2])))
    return (v / l).astype(f32)


def _cross(a, b):
    return _v(a[1] * b[2] - a[2] * b[1], a[2] * b[0] - a[0] * b[2], a[0] * b[1] - a[1] * b[0])


def f3x3_quat(c0, c1, c2):
    """math/quat_funcs.h:160-222 ([column][row] accessors)."""
    m00, m01, m02 = c0[0], c0[1], c0[2]
    m10, m11, m12 = c1[0], c1[1], c1[2]
    m20, m21, m22 = c2[0], c2[1], c2[2]
    x = f32(m00 - m11 - m22)
    y = f32(m11 - m00 - m22)
    z = f32(m22 - m00 - m11)
    w = f32(m00 + m11 + m22)
    b = max(max(x, y), max(z, w))
    i = 0
    if b == x:
        i = 1
    if b == y:
        i = 2
    if b == z:
        i = 3
    b = f32(np.sqrt(max(f32(b + f32(1.0)), f32(4.0) * f32(1.0 / 70368744177663.0))) * f32(0.5))
    mult = f32(f32(0.25) / b)
    if i == 0:
        r = _v((m12 - m21) * mult, (m20 - m02) * mult, (m01 - m10) * mult, b)
    elif i == 1:
        r = _v(b, (m01 + m10) * mult, (m20 + m02) * mult, (m12 - m21) * mult)
    elif i == 2:
        r = _v((m01 + m10) * mult, b, (m12 + m21) * mult, (m20 - m02) * mult)
    else:
        r = _v((m20 + m02) * mult, (m12 + m21) * mult, b, (m01 - m10) * mult)
    return r


def quat_lookat(forward, up):
    """math/quat_funcs.h:223-233."""
    forward = (-np.asarray(forward, f32)).astype(f32)
    right = _normalize3(_cross(np.asarray(up, f32), forward))
    up = _normalize3(_cross(forward, right))
    return f3x3_quat(right, up, forward)


def quat_f3x3(q):
    """math/quat_funcs.h:131-158; returns columns c0, c1, c2."""
    x, y, z, w = [f32(t) for t in q]
    xx, xy, xz = x * x, x * y, x * z
    yy, yz, zz = y * y, y * z, z * z
    wx, wy, wz = w * x, w * y, w * z
    one, two = f32(1.0), f32(2.0)
    c0 = _v(one - two * (yy + zz), two * (xy - wz), two * (xz + wy))
    c1 = _v(two * (xy + wz), one - two * (xx + zz), two * (yz - wx))
    c2 = _v(two * (xz - wy), two * (yz + wx), one - two * (xx + yy))
    return c0, c1, c2


def f4x4_trs(t, q, s):
    """math/float4x4_funcs.h:189-192: scale(rotate(translation(t), q), s), column major [16]."""
    c0, c1, c2 = quat_f3x3(q)
    # f4x4_mul_dir(translation, col) leaves xyz as is and w = 0
    M = np.zeros(16, f32)
    M[0:3] = c0 * f32(s[0])
    M[4:7] = c1 * f32(s[1])
    M[8:11] = c2 * f32(s[2])
    M[12:15] = np.asarray(t, f32)[:3]
    M[15] = 1.0
    return M


def quat_mul_dir(q, d):
    """math/quat_funcs.h:86-97."""
    qv = np.asarray(q[:3], f32)
    d = np.asarray(d, f32)
    uv = _cross(qv, d)
    uuv = _cross(qv, uv)
    y = (uv * f32(q[3])).astype(f32)
    y = (y + uuv).astype(f32)
    y = (y * f32(2.0)).astype(f32)
    return (y + d).astype(f32)


def srgb_inverse_eotf_fit(L):
    """math/color.h:212-218."""
    L = np.asarray(L, f32)
    l1 = np.sqrt(L).astype(f32)
    l2 = np.sqrt(l1).astype(f32)
    l3 = np.sqrt(l2).astype(f32)
    return (f32(0.658444) * l1 + f32(0.643378) * l2 - f32(0.298148) * l3).astype(f32)


def gamma_encode_rgba8(value):
    """GammaEncode_rgba8 = f4_rgba8(f4_sRGB_InverseEOTF_Fit(v)), math/color.h:102-112, 251-254."""
    v = srgb_inverse_eotf_fit(value)
    v = np.clip(v, f32(0.0), f32(1.0)).astype(f32)
    v = (v * f32(255.0) + f32(0.5)).astype(f32)
    b = v.astype(np.uint32)
    return np.array([b[0] | (b[1] << 8) | (b[2] << 16) | (b[3] << 24)], dtype=np.uint32)


def flat_texture(value):
    """GenFlatTexture, render_system.c:1003-1016."""
    return Texture(1, 1, gamma_encode_rgba8(np.asarray(value, f32)), TEX_RGBA8)


# ------------------------------------------------------------------- meshes

def gen_box_mesh():
    """GenBoxMesh, render_system.c:923-1001."""
    v = np.array([[1, 1, -1], [1, -1, -1], [1, 1, 1], [1, -1, 1],
                  [-1, 1, -1], [-1, -1, -1], [-1, 1, 1], [-1, -1, 1]], f32)
    vt = np.array([[0.875, 0.5], [0.625, 0.75], [0.625, 0.5], [0.375, 1.0], [0.375, 0.75],
                   [0.625, 0.0], [0.375, 0.25], [0.375, 0.0], [0.375, 0.5], [0.125, 0.75],
                   [0.125, 0.5], [0.625, 0.25], [0.875, 0.75], [0.625, 1.0]], f32)
    vn = np.array([[0, 1, 0], [0, 0, 1], [-1, 0, 0], [0, -1, 0], [1, 0, 0], [0, 0, -1]], f32)
    f = [(5, 1, 1), (3, 2, 1), (1, 3, 1), (3, 2, 2), (8, 4, 2), (4, 5, 2),
         (7, 6, 3), (6, 7, 3), (8, 8, 3), (2, 9, 4), (8, 10, 4), (6, 11, 4),
         (1, 3, 5), (4, 5, 5), (2, 9, 5), (5, 12, 6), (2, 9, 6), (6, 7, 6),
         (5, 1, 1), (7, 13, 1), (3, 2, 1), (3, 2, 2), (7, 14, 2), (8, 4, 2),
         (7, 6, 3), (5, 12, 3), (6, 7, 3), (2, 9, 4), (4, 5, 4), (8, 10, 4),
         (1, 3, 5), (3, 2, 5), (4, 5, 5), (5, 12, 6), (1, 3, 6), (2, 9, 6)]
    n = len(f)
    pos = np.zeros((n, 4), f32); nrm = np.zeros((n, 4), f32); uv = np.zeros((n, 4), f32)
    for i, (p, t, nn) in enumerate(f):
        pos[i, :3] = v[p - 1] * f32(0.5)
        uv[i, :2] = vt[t - 1]
        nrm[i, :3] = vn[nn - 1]
    return pos, nrm, uv


def gen_sphere_mesh():
    """GenSphereMesh, render_system.c:745-876 (sinf/cosf on float32 angles)."""
    vsteps, hsteps = 24, 48
    kPi, kTau = f32(3.1415926535897932), f32(6.283185307179586)
    dv, dh = f32(kPi / f32(vsteps)), f32(kTau / f32(hsteps))
    P, U = [], []
    sin, cos = lambda a: f32(np.sin(f32(a), dtype=f32)), lambda a: f32(np.cos(f32(a), dtype=f32))
    for vi in range(vsteps):
        th1, th2 = f32(f32(vi) * dv), f32(f32(vi + 1) * dv)
        st1, ct1, st2, ct2 = sin(th1), cos(th1), sin(th2), cos(th2)
        for h in range(hsteps):
            ph1, ph2 = f32(f32(h) * dh), f32(f32(h + 1) * dh)
            sp1, cp1, sp2, cp2 = sin(ph1), cos(ph1), sin(ph2), cos(ph2)
            one = f32(1.0)
            u1 = (f32(ph1 / kTau), f32(one - f32(th1 / kPi)))
            u2 = (f32(ph2 / kTau), f32(one - f32(th1 / kPi)))
            u3 = (f32(ph2 / kTau), f32(one - f32(th2 / kPi)))
            u4 = (f32(ph1 / kTau), f32(one - f32(th2 / kPi)))
            n1 = (st1 * cp1, ct1, st1 * sp1)
            n2 = (st1 * cp2, ct1, st1 * sp2)
            n3 = (st2 * cp2, ct2, st2 * sp2)
            n4 = (st2 * cp1, ct2, st2 * sp1)
            if vi == 0:
                P += [n1, n3, n4]; U += [u1, u3, u4]
            elif vi + 1 == vsteps:
                P += [n3, n1, n2]; U += [u3, u1, u2]
            else:
                P += [n1, n2, n4, n2, n3, n4]; U += [u1, u2, u4, u2, u3, u4]
    n = len(P)
    pos = np.zeros((n, 4), f32); uv = np.zeros((n, 4), f32)
    pos[:, :3] = np.array(P, f32)
    uv[:, :2] = np.array(U, f32)
    return pos, pos.copy(), uv


CUBE_FORWARDS = np.array([[1, 0, 0], [-1, 0, 0], [0, 1, 0], [0, -1, 0], [0, 0, 1], [0, 0, -1]], f32)
CUBE_UPS = np.array([[0, 1, 0], [0, 1, 0], [0, 0, -1], [0, 0, -1], [0, 1, 0], [0, 1, 0]], f32)
XP, XM, YP, YM, ZP, ZM = range(6)


def _lerp(a, b, t):
    return f32(f32(a) + f32(f32(b) - f32(a)) * f32(t))


def cornell_box(variant="boxes", brdf_lut=None, seed=1):
    """CmdCornellBox, render_system.c:1109-1306. variant: 'boxes' (pt_test) or 'spheres'."""
    desc = SceneDesc(brdf_lut=brdf_lut, seed=seed)
    box = gen_box_mesh()
    sphere = gen_sphere_mesh() if variant == "spheres" else None

    def create(mesh, t, q, s, albedo, rome, flags=0, ior=1.0):
        M = f4x4_trs(np.asarray(t, f32), q, np.asarray(s, f32))
        ta = len(desc.textures); desc.textures.append(flat_texture(albedo))   # Color_SDRToScene is identity (Rec709 scene)
        tr = len(desc.textures); desc.textures.append(flat_texture(rome))
        if rome[3] > 0.0:
            flags |= MATFLAG_EMISSIVE
        desc.entities.append(Entity(mesh[0], mesh[1], mesh[2], M, ta, tr, -1, flags, (0, 0, 0, 0), ior, 0.0))

    wall = f32(5.0)
    kDeci = f32(0.1)
    wall_scale = _v(f32(2.0) * wall, f32(2.0) * wall, kDeci, 0.0)
    cHi = f32(0.9); cLo = f32(f32(1.0) - cHi)
    red = _v(cHi, cLo, cLo, 1.0); green = _v(cLo, cHi, cLo, 1.0)
    blue = _v(cLo, cLo, cHi, 1.0); white = _v(cHi, cHi, cHi, 1.0)
    plastic = _v(0.9, 1.0, 0.0, 0.0); metal = _v(0.1, 1.0, 1.0, 0.0); light = _v(0.9, 1.0, 0.0, 1.0)
    F, Up = CUBE_FORWARDS, CUBE_UPS
    create(box, F[YM] * wall, quat_lookat(F[YP], Up[YP]), wall_scale, white, plastic)
    create(box, F[YP] * wall, quat_lookat(F[YM], Up[YM]), wall_scale, white, plastic)
    create(box, F[YP] * f32(wall - kDeci * f32(2.0)), quat_lookat(F[YM], Up[YM]), _v(1.0, 1.0, kDeci, 0.0),
           _v(1.0, 1.0, 1.0, 1.0), light)
    create(box, F[XM] * wall, quat_lookat(F[XP], Up[XP]), wall_scale, green, plastic)
    create(box, F[XP] * wall, quat_lookat(F[XM], Up[XM]), wall_scale, red, plastic)
    create(box, F[ZP] * wall, quat_lookat(F[ZP], Up[ZP]), wall_scale, white, plastic)
    create(box, F[ZM] * wall, quat_lookat(F[ZM], Up[ZM]), wall_scale, blue, plastic)
    if variant == "spheres":
        sphere_scale = f32(0.75); margin = f32(sphere_scale * f32(1.5))
        lo = f32(-wall + margin); hi = f32(wall - margin)
        mat_flags = [0, 0, MATFLAG_REFRACTIVE]
        metallics = [1.0, 0.0, 0.0]
        iors = [1.0, 1.0, 1.5]
        quat_id = _v(0, 0, 0, 1)
        for row in range(3):
            z = _lerp(lo, hi, f32(f32(row + 0.5) / f32(3)))
            for col in range(5):
                tcol = f32(f32(col + 0.5) / f32(5))
                rough = _lerp(0.0, 1.0, tcol)
                x = _lerp(lo, hi, tcol)
                create(sphere, _v(x, lo, z, 0.0), quat_id, _v(sphere_scale, sphere_scale, sphere_scale, sphere_scale),
                       white, _v(rough, 1.0, metallics[row], 0.0), mat_flags[row], iors[row])
    else:
        box_scale = f32(2.0); margin = f32(box_scale * f32(0.5))
        lo = f32(-wall + margin); hi = f32(wall - margin)
        create(box, _v(_lerp(lo, hi, 0.2), f32(-wall + box_scale), _lerp(lo, hi, 0.2), 0.0),
               quat_lookat(_normalize3(_v(0.2, 0.0, 1.0)), F[YP]),
               _v(box_scale, box_scale * f32(2.0), box_scale, 0.0), white, metal)
        create(box, _v(_lerp(lo, hi, 0.8), f32(-wall + box_scale * f32(0.5)), _lerp(lo, hi, 0.8), 0.0),
               quat_lookat(_normalize3(_v(-0.2, 0.0, 1.0)), F[YP]),
               _v(box_scale, box_scale, box_scale, 0.0), white, plastic)
    return desc


def textured_room(brdf_lut=None, seed=1, tex_seed=5, features=("albedo", "rome", "normal", "emission", "sky")):
    """Cornell boxes with every material input of GetSurface in play (a parity scene, not one of the
    reference's): varying RGBA8 albedo textures (true bilinear + gamma decode), ROME textures with varying
    roughness / metallic and an emission channel that is only partly on (EmissionPdf's texel probes), RG16
    normal maps, a sky-flagged ceiling and a sky cubemap behind it."""
    desc = cornell_box("boxes", brdf_lut=brdf_lut, seed=seed)
    g = np.random.Generator(np.random.PCG64(tex_seed))

    def rgba8(r, gc, b, a):
        q = lambda x: np.clip(np.rint(np.asarray(x, np.float64) * 255.0), 0, 255).astype(np.uint32)
        return (q(r) | (q(gc) << 8) | (q(b) << 16) | (q(a) << 24)).astype(np.uint32).reshape(-1)

    def add(tex):
        desc.textures.append(tex)
        return len(desc.textures) - 1

    def albedo_tex(n, tint):
        v = g.uniform(0.35, 0.95, (n, n, 3)) * np.asarray(tint)
        return add(Texture(n, n, rgba8(v[..., 0], v[..., 1], v[..., 2], np.ones((n, n))), TEX_RGBA8))

    def rome_tex(n, metal_frac, emit=None):
        rough = g.uniform(0.05, 1.0, (n, n))
        metal = (g.uniform(0, 1, (n, n)) < metal_frac).astype(np.float64)
        em = np.zeros((n, n)) if emit is None else emit
        return add(Texture(n, n, rgba8(rough, np.ones((n, n)), metal, em), TEX_RGBA8))

    def normal_tex(n, tilt):
        xy = g.uniform(-tilt, tilt, (n, n, 2))
        q = np.clip(np.rint(xy * 32767.0), -32767, 32767).astype(np.int32)
        packed = (q[..., 0] & 0xFFFF).astype(np.uint32) | ((q[..., 1] & 0xFFFF).astype(np.uint32) << 16)
        return add(Texture(n, n, packed.reshape(-1), TEX_RG16))

    e = desc.entities
    # 0 floor, 1 ceiling, 2 light, 3 -X wall, 4 +X wall, 5 +Z wall, 6 -Z wall, 7 tall box, 8 small box
    if "albedo" in features:
        e[0].albedo, e[3].albedo = albedo_tex(32, (1, 1, 1)), albedo_tex(32, (0.3, 1, 0.3))
        e[4].albedo, e[5].albedo = albedo_tex(16, (1, 0.3, 0.3)), albedo_tex(64, (1, 1, 1))
        e[7].albedo = albedo_tex(8, (1, 0.9, 0.6))
    if "rome" in features:
        e[0].rome, e[3].rome, e[4].rome = rome_tex(16, 0.0), rome_tex(8, 0.5), rome_tex(8, 0.0)
        e[7].rome, e[8].rome = rome_tex(4, 1.0), rome_tex(16, 0.3)
    if "normal" in features:
        e[0].normal, e[4].normal, e[7].normal = normal_tex(16, 0.35), normal_tex(32, 0.5), normal_tex(8, 0.2)
    if "emission" in features:      # the light: emission on a checkerboard of its texture
        yy, xx = np.mgrid[0:8, 0:8]
        e[2].rome = rome_tex(8, 0.0, emit=((xx + yy) % 2).astype(np.float64))
        e[2].flags |= MATFLAG_EMISSIVE
    if "sky" in features:           # the ceiling is a window to the sky
        e[1].flags |= MATFLAG_SKY
    n = 16
    yy, xx = np.mgrid[0:n, 0:n]
    sky = np.zeros((6, n, n, 3), f32)
    for f in range(6):
        sky[f, ..., 0] = 0.2 + 0.05 * f + 0.3 * xx / n
        sky[f, ..., 1] = 0.3 + 0.4 * yy / n
        sky[f, ..., 2] = 0.8 - 0.03 * f
    sky[2] *= 3.0       # brighter straight up
    desc.sky = sky
    return desc


# ------------------------------------------------------------ synthetic inputs

class Pcg32:
    """O'Neill's minimal pcg32 (the generator BASELINE.json names for inputs), vectorised."""

    def __init__(self, seed, stream=54):
        self.inc = np.uint64((stream << 1) | 1)
        self.state = np.uint64(0)
        self._step()
        self.state = np.uint64((int(self.state) + seed) & 0xFFFFFFFFFFFFFFFF)
        self._step()

    def _step(self):
        self.state = np.uint64((int(self.state) * 6364136223846793005 + int(self.inc)) & 0xFFFFFFFFFFFFFFFF)

    def next_u32(self):
        old = int(self.state)
        self._step()
        xorshifted = (((old >> 18) ^ old) >> 27) & 0xFFFFFFFF
        rot = old >> 59
        return ((xorshifted >> rot) | (xorshifted << ((-rot) & 31))) & 0xFFFFFFFF

    def generator(self):
        """numpy Generator seeded from this pcg32 stream (bulk draws)."""
        return np.random.Generator(np.random.PCG64(self.next_u32() | (self.next_u32() << 32)))


def random_soup(tri_count, seed=11, extent=100.0, edge_lo=0.02, edge_hi=0.5):
    """C5 mesh: centroids uniform in [0,extent)^3, edge length log-uniform, random orientation."""
    g = Pcg32(seed).generator()
    c = g.uniform(0.0, extent, (tri_count, 1, 3))
    edge = np.exp(g.uniform(np.log(edge_lo), np.log(edge_hi), (tri_count, 1, 1)))
    d = g.normal(size=(tri_count, 3, 3))
    d /= np.linalg.norm(d, axis=2, keepdims=True)
    v = (c + d * edge * 0.5).astype(f32)
    pos = np.zeros((tri_count * 3, 4), f32)
    pos[:, :3] = v.reshape(-1, 3)
    pos[:, 3] = 1.0
    n = np.cross(v[:, 1] - v[:, 0], v[:, 2] - v[:, 0])
    n /= np.maximum(np.linalg.norm(n, axis=1, keepdims=True), 1e-20)
    nrm = np.zeros_like(pos)
    nrm[:, :3] = np.repeat(n.astype(f32), 3, axis=0)
    uv = np.zeros_like(pos)
    desc = SceneDesc(seed=1)
    desc.textures.append(flat_texture(_v(0.5, 0.5, 0.5, 1.0)))
    desc.textures.append(flat_texture(_v(0.9, 1.0, 0.0, 0.0)))
    M = np.eye(4, dtype=f32).reshape(-1)
    desc.entities.append(Entity(pos, nrm, uv, M, 0, 1, -1, 0))
    return desc


def random_rays(n, seed=13, extent=100.0, t_far=8388608.0, occlusion=False):
    """C5 rays: origins uniform in the box, directions uniform on the sphere."""
    g = Pcg32(seed).generator()
    ro = np.zeros((n, 4), f32)
    ro[:, :3] = g.uniform(0.0, extent, (n, 3))
    phi = g.uniform(0.0, 2 * np.pi, n)
    ct = g.uniform(-1.0, 1.0, n)
    stt = np.sqrt(np.maximum(1.0 - ct * ct, 0.0))
    rd = np.zeros((n, 4), f32)
    rd[:, 0] = stt * np.cos(phi); rd[:, 1] = stt * np.sin(phi); rd[:, 2] = ct
    rd[:, 3] = g.uniform(1.0, 50.0, n) if occlusion else t_far
    return ro, rd


def interior(target_tris=1_000_000, seed=7, rooms=6, emissive_fraction=0.005):
    """C3 stand-in: a grid of rooms with tessellated, displaced walls and a few emissive panels.

    The reference loads such a scene through cgltf (src/rendering/gltf_model.c); here the
    same flattened input (unindexed triangles, per-entity material, RGBA8 textures with
    emission in ROME alpha) is generated directly.
    """
    g = Pcg32(seed).generator()
    desc = SceneDesc(seed=1)
    room = 6.0
    quads_per_room = 6
    tess = max(2, int(np.sqrt(target_tris / (rooms * rooms * quads_per_room * 2.0))))
    tex_albedo = [len(desc.textures) + i for i in range(4)]
    for i in range(4):
        col = g.uniform(0.3, 0.9, 3)
        tex = np.zeros((256, 256), np.uint32)
        base = gamma_encode_rgba8(_v(col[0], col[1], col[2], 1.0))[0]
        noise = g.integers(0, 12, (256, 256)).astype(np.uint32)
        tex[:] = base
        tex = (tex & np.uint32(0xFFFFFF00)) | np.minimum((tex & np.uint32(0xFF)) + noise, 255).astype(np.uint32)
        desc.textures.append(Texture(256, 256, tex.reshape(-1), TEX_RGBA8))
    rome_plain = len(desc.textures); desc.textures.append(flat_texture(_v(0.8, 1.0, 0.0, 0.0)))
    rome_light = len(desc.textures); desc.textures.append(flat_texture(_v(0.9, 1.0, 0.0, 1.0)))
    white = len(desc.textures); desc.textures.append(flat_texture(_v(1.0, 1.0, 1.0, 1.0)))

    def quad(origin, du, dv, n_u, n_v, amp):
        u = np.linspace(0, 1, n_u + 1); v = np.linspace(0, 1, n_v + 1)
        U, V = np.meshgrid(u, v, indexing="ij")
        normal = np.cross(du, dv); normal /= np.linalg.norm(normal)
        disp = g.uniform(-amp, amp, U.shape)
        disp[0, :] = disp[-1, :] = 0; disp[:, 0] = disp[:, -1] = 0
        P = origin + U[..., None] * du + V[..., None] * dv + disp[..., None] * normal
        a = P[:-1, :-1]; b = P[1:, :-1]; c = P[1:, 1:]; d = P[:-1, 1:]
        tris = np.stack([a, b, c, a, c, d], axis=2).reshape(-1, 3)
        UV = np.stack([U, V], axis=-1)
        ua = UV[:-1, :-1]; ub = UV[1:, :-1]; uc = UV[1:, 1:]; ud = UV[:-1, 1:]
        uvs = np.stack([ua, ub, uc, ua, uc, ud], axis=2).reshape(-1, 2)
        return tris, uvs, normal

    for rx in range(rooms):
        for rz in range(rooms):
            o = np.array([rx * room, 0.0, rz * room])
            X, Y, Z = np.array([room, 0, 0.0]), np.array([0, room * 0.5, 0.0]), np.array([0, 0, room])
            # inward-facing quads: floor, ceiling, 4 walls (walls have door gaps via shorter span)
            faces = [(o, Z, X), (o + Y, X, Z), (o, X * 0.8, Y), (o + Z, Y, X * 0.8), (o, Y, Z * 0.8), (o + X, Z * 0.8, Y)]
            for k, (org, du, dv) in enumerate(faces):
                tris, uvs, normal = quad(org, du, dv, tess, tess, 0.01)
                pos = np.zeros((len(tris), 4), f32); pos[:, :3] = tris; pos[:, 3] = 1
                nr = np.zeros_like(pos); nr[:, :3] = normal
                uv = np.zeros_like(pos); uv[:, :2] = uvs * 4.0
                M = np.eye(4, dtype=f32).reshape(-1)
                desc.entities.append(Entity(pos, nr, uv, M, tex_albedo[(rx + rz + k) % 4], rome_plain, -1, 0))
            # one emissive ceiling panel per room, tessellated so ~emissive_fraction of all triangles emit
            pt = max(1, int(tess * np.sqrt(emissive_fraction * quads_per_room)))
            tris, uvs, normal = quad(o + Y * 0.98 + X * 0.35 + Z * 0.35, X * 0.3, Z * 0.3, pt, pt, 0.0)
            pos = np.zeros((len(tris), 4), f32); pos[:, :3] = tris; pos[:, 3] = 1
            nr = np.zeros_like(pos); nr[:, :3] = normal
            uv = np.zeros_like(pos); uv[:, :2] = uvs
            desc.entities.append(Entity(pos, nr, uv, np.eye(4, dtype=f32).reshape(-1), white, rome_light, -1, MATFLAG_EMISSIVE))
    return desc


def camera_lookat(eye, at, fovy=90.0):
    """CmdTeleport + CmdLookat, render_system.c:1308-1346. Returns (position[4], rotation[4], fovy)."""
    eye = np.asarray(list(eye) + [0.0] * (4 - len(eye)), f32)
    at = np.asarray(list(at) + [0.0] * (4 - len(at)), f32)
    rd = _normalize3((at - eye)[:3])
    rot = quat_lookat(rd, _v(0.0, 1.0, 0.0))
    return eye, rot, float(fovy)
