"""Host-side set-up math of the `cudapath` plugin: everything the reference computes once per
scene in plugin constructors / configure() and that the flattened scene description carries as
plain numbers (transforms, camera matrices, filter tables, derived BSDF constants, LUT slices).

Not on the hot path.  Citations are reference paths (relative to the reference root).
"""
import math
import os

import numpy as np

F = np.float32

# Named indices of refraction, src/bsdfs/ior.h:39-64
IOR_TABLE = {
    "vacuum": 1.0, "helium": 1.000036, "hydrogen": 1.000132, "air": 1.000277, "carbon dioxide": 1.00045,
    "water": 1.3330, "acetone": 1.36, "ethanol": 1.361, "carbon tetrachloride": 1.461, "glycerol": 1.4729,
    "benzene": 1.501, "silicone oil": 1.52045, "bromine": 1.661, "water ice": 1.31, "fused quartz": 1.458,
    "pyrex": 1.470, "acrylic glass": 1.49, "polypropylene": 1.49, "bk7": 1.5046, "sodium chloride": 1.544,
    "amber": 1.55, "pet": 1.5750, "diamond": 2.419,
}


def lookup_ior(value, default):
    """lookupIOR(props, name, default), src/bsdfs/ior.h:95-111."""
    if value is None:
        value = default
    if isinstance(value, (int, float)):
        return float(F(value))
    key = str(value).lower()
    if key not in IOR_TABLE:
        raise ValueError('Unable to find an IOR value for "%s"! Valid choices are: %s' % (key, ", ".join(IOR_TABLE)))
    return float(F(IOR_TABLE[key]))


# ---------------------------------------------------------------------------
# Transform (src/libcore/transform.cpp): row-major 4x4, column vectors
# ---------------------------------------------------------------------------
def identity():
    return np.eye(4, dtype=np.float64)


def translate(v):
    m = identity()
    m[:3, 3] = v
    return m


def scale(v):
    v = np.broadcast_to(np.asarray(v, dtype=np.float64), (3,))
    return np.diag([v[0], v[1], v[2], 1.0])


def rotate(axis, angle_deg):
    """Transform::rotate(axis, angle) (transform.cpp:60-96)."""
    a = np.asarray(axis, dtype=np.float64)
    a = a / np.linalg.norm(a)
    s, c = math.sin(math.radians(angle_deg)), math.cos(math.radians(angle_deg))
    x, y, z = a
    m = identity()
    m[0, 0] = x * x + (1 - x * x) * c
    m[0, 1] = x * y * (1 - c) - z * s
    m[0, 2] = x * z * (1 - c) + y * s
    m[1, 0] = x * y * (1 - c) + z * s
    m[1, 1] = y * y + (1 - y * y) * c
    m[1, 2] = y * z * (1 - c) - x * s
    m[2, 0] = x * z * (1 - c) - y * s
    m[2, 1] = y * z * (1 - c) + x * s
    m[2, 2] = z * z + (1 - z * z) * c
    return m


def look_at(origin, target, up):
    """Transform::lookAt (transform.cpp:168-200): camera-to-world with +z forward, and the
    `left = cross(up, dir)` handedness Mitsuba uses."""
    p, t, u = (np.asarray(v, dtype=np.float64) for v in (origin, target, up))
    d = t - p
    d = d / np.linalg.norm(d)
    left = np.cross(u / np.linalg.norm(u), d)
    left = left / np.linalg.norm(left)
    newup = np.cross(d, left)
    m = identity()
    m[:3, 0], m[:3, 1], m[:3, 2], m[:3, 3] = left, newup, d, p
    return m


def perspective(fov_deg, near, far):
    """Transform::perspective (transform.cpp:99-121)."""
    recip = 1.0 / (far - near)
    cot = 1.0 / math.tan(math.radians(fov_deg / 2.0))
    return np.array([[cot, 0, 0, 0], [0, cot, 0, 0], [0, 0, far * recip, -near * far * recip], [0, 0, 1, 0]],
                    dtype=np.float64)


def xform_point(m, p):
    q = m @ np.array([p[0], p[1], p[2], 1.0])
    return q[:3] / q[3]


def xform_vector(m, v):
    return m[:3, :3] @ np.asarray(v, dtype=np.float64)


def xform_normal(m, n):
    return np.linalg.inv(m)[:3, :3].T @ np.asarray(n, dtype=np.float64)


# ---------------------------------------------------------------------------
# Sensor (src/librender/sensor.cpp:84-109,239-310; src/sensors/perspective.cpp:126-178)
# ---------------------------------------------------------------------------
def x_fov(fov, fov_axis, aspect):
    axis = fov_axis.lower()
    if axis == "smaller":
        axis = "y" if aspect > 1 else "x"
    elif axis == "larger":
        axis = "x" if aspect > 1 else "y"
    if axis == "x":
        return fov
    if axis == "y":
        return math.degrees(2 * math.atan(math.tan(0.5 * math.radians(fov)) * aspect))
    if axis == "diagonal":
        diagonal = 2 * math.tan(0.5 * math.radians(fov))
        width = diagonal / math.sqrt(1.0 + 1.0 / (aspect * aspect))
        return math.degrees(2 * math.atan(width * 0.5))
    raise ValueError("The 'fovAxis' parameter must be set to one of 'smaller', 'larger', 'diagonal', 'x', or 'y'!")


def camera_matrices(to_world, xfov, near, far, film_size, crop_size=None, crop_offset=(0, 0)):
    """m_cameraToSample / m_sampleToCamera / m_dx / m_dy of PerspectiveCameraImpl::configure."""
    if not (0 < xfov < 180):
        raise ValueError("The horizontal field of view must be in the interval (0, 180)!")
    fw, fh = film_size
    cw, ch = crop_size if crop_size is not None else film_size
    aspect = fw / float(fh)
    rel_size = (cw / float(fw), ch / float(fh))
    rel_off = (crop_offset[0] / float(fw), crop_offset[1] / float(fh))
    camera_to_sample = (scale([1.0 / rel_size[0], 1.0 / rel_size[1], 1.0])
                        @ translate([-rel_off[0], -rel_off[1], 0.0])
                        @ scale([-0.5, -0.5 * aspect, 1.0])
                        @ translate([-1.0, -1.0 / aspect, 0.0])
                        @ perspective(xfov, near, far))
    s2c = np.linalg.inv(camera_to_sample)
    inv_res = (1.0 / cw, 1.0 / ch)
    p0 = xform_point(s2c, [0, 0, 0])
    dx = xform_point(s2c, [inv_res[0], 0, 0]) - p0
    dy = xform_point(s2c, [0, inv_res[1], 0]) - p0
    return s2c, dx, dy, inv_res


# ---------------------------------------------------------------------------
# Reconstruction filters (src/libcore/rfilter.cpp:37-56, src/rfilters/{box,gaussian}.cpp)
# ---------------------------------------------------------------------------
FILTER_RESOLUTION = 31


def filter_table(kind="gaussian", radius=None, stddev=0.5):
    kind = kind.lower()
    if kind == "box":
        r = F((0.5 if radius is None else radius)) + F(1e-5)

        def ev(x):
            return F(1.0) if abs(x) <= r else F(0.0)
    elif kind == "gaussian":
        sd = F(stddev)
        r = F(4) * sd
        alpha = F(-1.0) / (F(2.0) * sd * sd)

        def ev(x):
            # math::fastexp goes through double on Linux (include/mitsuba/core/math.h:176-199)
            a = F(math.exp(float(alpha * x * x)))
            b = F(math.exp(float(alpha * r * r)))
            return max(F(0.0), F(a - b))
    else:
        raise ValueError("unsupported reconstruction filter '%s' (in scope: box, gaussian)" % kind)
    vals = np.zeros(FILTER_RESOLUTION + 1, dtype=F)
    s = F(0)
    for i in range(FILTER_RESOLUTION):
        v = ev(F(F(r * F(i)) / F(FILTER_RESOLUTION)))
        vals[i] = v
        s = F(s + v)
    s = F(s * F(F(2) * r / F(FILTER_RESOLUTION)))
    norm = F(1.0) / s
    vals[:FILTER_RESOLUTION] *= norm
    return float(r), vals


# ---------------------------------------------------------------------------
# Fresnel helpers (src/libcore/util.cpp:651-681, 814-860) and Gauss-Lobatto (src/libcore/quad.cpp:287-411)
# ---------------------------------------------------------------------------
def fresnel_dielectric_ext(cos_i, eta):
    cos_i, eta = F(cos_i), F(eta)
    if eta == 1:
        return F(0.0)
    sc = F(1) / eta if cos_i > 0 else eta
    cos_t_sqr = F(1) - (F(1) - cos_i * cos_i) * (sc * sc)
    if cos_t_sqr <= 0:
        return F(1.0)
    ci, ct = abs(cos_i), F(np.sqrt(cos_t_sqr))
    rs = (ci - eta * ct) / (ci + eta * ct)
    rp = (eta * ci - ct) / (eta * ci + ct)
    return F(0.5) * (rs * rs + rp * rp)


_GL_ALPHA, _GL_BETA = F(math.sqrt(2.0 / 3.0)), F(1.0 / math.sqrt(5.0))
_GL_X1, _GL_X2, _GL_X3 = F(0.94288241569547971906), F(0.64185334234578130578), F(0.23638319966214988028)


def gauss_lobatto(f, a, b, max_evals=1024, abs_err=0.0, rel_err=1e-5):
    """GaussLobattoIntegrator(maxEvals, absError, relError).integrate(f, a, b) in fp32."""
    a, b = F(a), F(b)
    evals = [0]
    eps = np.finfo(F).eps

    m, h = (a + b) / F(2), (b - a) / F(2)
    y1, y3, y5, y7 = f(a), f(m - _GL_ALPHA * h), f(m - _GL_BETA * h), f(m)
    y9, y11, y13 = f(m + _GL_BETA * h), f(m + _GL_ALPHA * h), f(b)
    acc = h * (F(0.0158271919734801831) * (y1 + y13)
               + F(0.0942738402188500455) * (f(m - _GL_X1 * h) + f(m + _GL_X1 * h))
               + F(0.1550719873365853963) * (y3 + y11)
               + F(0.1888215739601824544) * (f(m - _GL_X2 * h) + f(m + _GL_X2 * h))
               + F(0.1997734052268585268) * (y5 + y9)
               + F(0.2249264653333395270) * (f(m - _GL_X3 * h) + f(m + _GL_X3 * h))
               + F(0.2426110719014077338) * y7)
    evals[0] += 13
    integral2 = (h / F(6)) * (y1 + y13 + F(5) * (y5 + y9))
    integral1 = (h / F(1470)) * (F(77) * (y1 + y13) + F(432) * (y3 + y11) + F(625) * (y5 + y9) + F(672) * y7)
    r = F(1.0)
    if abs(integral2 - acc) != 0.0:
        r = abs(integral1 - acc) / abs(integral2 - acc)
    if r == 0.0 or r > 1.0:
        r = F(1.0)
    tol = F(np.inf)
    if rel_err != 0 and acc != 0:
        tol = acc * max(F(rel_err), eps) / (r * eps)
    if abs_err != 0:
        tol = min(tol, F(abs_err) / (r * eps))
    evals[0] += 2

    def step(a, b, fa, fb):
        h, m = (b - a) / F(2), (a + b) / F(2)
        mll, ml, mr, mrr = m - _GL_ALPHA * h, m - _GL_BETA * h, m + _GL_BETA * h, m + _GL_ALPHA * h
        fmll, fml, fm, fmr, fmrr = f(mll), f(ml), f(m), f(mr), f(mrr)
        i2 = (h / F(6)) * (fa + fb + F(5) * (fml + fmr))
        i1 = (h / F(1470)) * (F(77) * (fa + fb) + F(432) * (fmll + fmrr) + F(625) * (fml + fmr) + F(672) * fm)
        evals[0] += 5
        if evals[0] >= max_evals:
            return i1
        dist = F(tol + (i1 - i2))
        if dist == tol or mll <= a or b <= mrr:
            return i1
        return (step(a, mll, fa, fmll) + step(mll, ml, fmll, fml) + step(ml, m, fml, fm)
                + step(m, mr, fm, fmr) + step(mr, mrr, fmr, fmrr) + step(mrr, b, fmrr, fb))

    with np.errstate(over="ignore", invalid="ignore"):
        return float(step(a, b, f(a), f(b)))


def fresnel_diffuse_reflectance(eta):
    """fresnelDiffuseReflectance(eta, fast=false), src/libcore/util.cpp:856-860."""
    return gauss_lobatto(lambda xi: fresnel_dielectric_ext(F(np.sqrt(F(xi))), eta), 0.0, 1.0, 1024, 0.0, 1e-5)


# ---------------------------------------------------------------------------
# Rough transmittance LUT (src/bsdfs/rtrans.h) + cubic interpolation (src/libcore/spline.cpp)
# ---------------------------------------------------------------------------
def _cubic_weights(p, size):
    t = F(F(p) * F(size - 1))
    knot = min(int(t), size - 2)
    t = F(t - F(knot))
    t2, t3 = t * t, t * t * t
    w = [F(0), F(2) * t3 - F(3) * t2 + F(1), F(-2) * t3 + F(3) * t2, F(0)]
    d0, d1 = t3 - F(2) * t2 + t, t3 - t2
    if knot > 0:
        w[2] += F(0.5) * d0
        w[0] -= F(0.5) * d0
    else:
        w[2] += d0
        w[1] -= d0
    if knot + 2 < size:
        w[3] += F(0.5) * d1
        w[1] -= F(0.5) * d1
    else:
        w[2] += d1
        w[1] -= d1
    return knot, w


def cubic_interp_nd(p, values):
    """evalCubicInterp{1,2,3}D with min=0, max=1.  `values` is indexed [..., dim1, dim0];
    p = (p0, p1, ...) with p0 the fastest-varying dimension (spline.cpp:236-297,379-450).
    Note: the 1-D reference routine uses explicit derivative estimates; it is algebraically the
    same stencil, and cubic_interp_1d below restates it literally."""
    nd = len(p)
    shape = values.shape
    knots, weights = [], []
    for dim in range(nd):
        if not (0.0 <= p[dim] <= 1.0):
            return F(0.0)
        k, w = _cubic_weights(p[dim], shape[nd - 1 - dim])
        knots.append(k)
        weights.append(w)
    result = F(0.0)
    if nd == 2:
        for y in range(-1, 3):
            wy = weights[1][y + 1]
            for x in range(-1, 3):
                wxy = F(weights[0][x + 1] * wy)
                if wxy == 0:
                    continue
                result = F(result + values[knots[1] + y, knots[0] + x] * wxy)
    elif nd == 3:
        for z in range(-1, 3):
            wz = weights[2][z + 1]
            for y in range(-1, 3):
                wyz = F(weights[1][y + 1] * wz)
                for x in range(-1, 3):
                    wxyz = F(weights[0][x + 1] * wyz)
                    if wxyz == 0:
                        continue
                    result = F(result + values[knots[2] + z, knots[1] + y, knots[0] + x] * wxyz)
    else:
        raise ValueError(nd)
    return result


def cubic_interp_1d(x, values):
    """evalCubicInterp1D(x, values, size, 0, 1) (spline.cpp:23-58)."""
    size = len(values)
    if not (0.0 <= x <= 1.0):
        return F(0.0)
    t = F(F(x) * F(size - 1))
    k = max(0, min(int(t), size - 2))
    f0, f1 = values[k], values[k + 1]
    d0 = F(0.5) * (values[k + 1] - values[k - 1]) if k > 0 else values[k + 1] - values[k]
    d1 = F(0.5) * (values[k + 2] - values[k]) if k + 2 < size else values[k + 1] - values[k]
    t = F(t - F(k))
    t2, t3 = t * t, t * t * t
    return F((F(2) * t3 - F(3) * t2 + F(1)) * f0 + (F(-2) * t3 + F(3) * t2) * f1 + (t3 - F(2) * t2 + t) * d0 + (t3 - t2) * d1)


def data_dir():
    """Where data/microfacet/*.dat and data/ior/*.spd live: $MTS_DATA_DIR, the package's own
    data directory, or the reference checkout when it is mounted (this container only)."""
    cands = [os.environ.get("MTS_DATA_DIR"), os.path.join(os.path.dirname(os.path.abspath(__file__)), "data"),
             os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle", "_ref", "data"),
             "/root/reference/data"]
    for c in cands:
        if c and os.path.isdir(os.path.join(c, "microfacet")):
            return c
    raise FileNotFoundError("Mitsuba data directory (data/microfacet/*.dat) not found; set MTS_DATA_DIR")


class RoughTransmittance:
    """RoughTransmittance file loader + setEta/setAlpha reductions (rtrans.h:72-150,292-390)."""

    _cache = {}

    def __init__(self, kind):
        if kind not in self._cache:
            path = os.path.join(data_dir(), "microfacet", "%s.dat" % kind)
            raw = open(path, "rb").read()
            header = b"MTS_TRANSMITTANCE"
            if raw[:len(header)] != header:
                raise ValueError("Encountered an invalid transmittance data file!")
            off = len(header)
            n_eta, n_alpha, n_theta = (int(v) for v in np.frombuffer(raw, dtype="<u8", count=3, offset=off))
            off += 24
            rng = np.frombuffer(raw, dtype="<f4", count=4, offset=off)
            off += 16
            total = 2 * n_eta * n_alpha * (n_theta + 1)
            data = np.frombuffer(raw, dtype="<f4", count=total, offset=off).reshape(2 * n_eta, n_alpha, n_theta + 1)
            assert off + 4 * total == len(raw)
            self._cache[kind] = (n_eta, n_alpha, n_theta, tuple(float(v) for v in rng),
                                 np.ascontiguousarray(data[:, :, :n_theta]), np.ascontiguousarray(data[:, :, n_theta]))
        (self.n_eta, self.n_alpha, self.n_theta, (self.eta_min, self.eta_max, self.alpha_min, self.alpha_max),
         self.trans, self.diff_trans) = self._cache[kind]

    def check_eta(self, eta):
        e = 1 / eta if eta < 1 else eta
        if e < self.eta_min or e > self.eta_max:
            raise ValueError("Error: the requested relative index of refraction eta=%f is outside of the "
                             "supported range [%f, %f]!" % (e, self.eta_min, self.eta_max))

    def check_alpha(self, alpha):
        if alpha < self.alpha_min or alpha > self.alpha_max:
            raise ValueError("Error: the requested roughness value alpha=%f is outside of the supported range "
                             "[%f, %f]!" % (alpha, self.alpha_min, self.alpha_max))

    def reduce_eta(self, eta):
        """setEta: -> (trans[nAlpha, nTheta], diffTrans[nAlpha])."""
        eta = F(eta)
        trans, diff = self.trans[:self.n_eta], self.diff_trans[:self.n_eta]
        if eta < 1:
            trans, diff = self.trans[self.n_eta:], self.diff_trans[self.n_eta:]
            eta = F(1.0) / eta
        if eta < self.eta_min:
            eta = F(self.eta_min)
        warped_eta = F(np.power(F((eta - F(self.eta_min)) / F(self.eta_max - self.eta_min)), F(0.25)))
        d_alpha, d_theta = F(1.0) / F(self.n_alpha - 1), F(1.0) / F(self.n_theta - 1)
        new_trans = np.zeros((self.n_alpha, self.n_theta), dtype=F)
        new_diff = np.zeros((self.n_alpha,), dtype=F)
        for i in range(self.n_alpha):
            for j in range(self.n_theta):
                new_trans[i, j] = cubic_interp_nd((F(j) * d_theta, F(i) * d_alpha, warped_eta), trans)
            new_diff[i] = cubic_interp_nd((F(i) * d_alpha, warped_eta), diff)
        return new_trans, new_diff

    def warp_alpha(self, alpha):
        return F(np.power(F((F(alpha) - F(self.alpha_min)) / F(self.alpha_max - self.alpha_min)), F(0.25)))

    def reduce_alpha(self, trans2d, diff1d, alpha):
        """setAlpha: -> (trans[nTheta], diffTrans scalar)."""
        wa = self.warp_alpha(alpha)
        d_theta = F(1.0) / F(self.n_theta - 1)
        out = np.array([cubic_interp_nd((F(i) * d_theta, wa), trans2d) for i in range(self.n_theta)], dtype=F)
        return out, cubic_interp_1d(wa, diff1d)


# ---------------------------------------------------------------------------
# Mesh helpers
# ---------------------------------------------------------------------------
def uv_tangents(positions, texcoords, indices):
    """TriMesh::computeUVTangents (src/librender/trimesh.cpp:683-743), fp32."""
    p = np.asarray(positions, dtype=F)
    uv = np.asarray(texcoords, dtype=F)
    idx = np.asarray(indices, dtype=np.int64).reshape(-1, 3)
    v0, v1, v2 = p[idx[:, 0]], p[idx[:, 1]], p[idx[:, 2]]
    t0, t1, t2 = uv[idx[:, 0]], uv[idx[:, 1]], uv[idx[:, 2]]
    dp1, dp2 = v1 - v0, v2 - v0
    duv1, duv2 = t1 - t0, t2 - t0
    n = np.cross(dp1, dp2)
    length = np.linalg.norm(n, axis=1).astype(F)
    det = (duv1[:, 0] * duv2[:, 1] - duv1[:, 1] * duv2[:, 0]).astype(F)
    out = np.zeros((idx.shape[0], 6), dtype=F)
    ok = (length != 0) & (det != 0)
    with np.errstate(divide="ignore", invalid="ignore"):
        inv = (F(1.0) / det)[:, None]
        dpdu = (duv2[:, 1:2] * dp1 - duv1[:, 1:2] * dp2) * inv
        dpdv = (-duv2[:, 0:1] * dp1 + duv1[:, 0:1] * dp2) * inv
    out[ok, 0:3] = dpdu[ok]
    out[ok, 3:6] = dpdv[ok]
    deg = (length != 0) & (det == 0)
    for i in np.nonzero(deg)[0]:  # coordinateSystem(n/length, dpdu, dpdv) (util.cpp:592-601)
        a = n[i] / length[i]
        if abs(a[0]) > abs(a[1]):
            il = F(1.0) / F(np.sqrt(a[0] * a[0] + a[2] * a[2]))
            c = np.array([a[2] * il, 0.0, -a[0] * il], dtype=F)
        else:
            il = F(1.0) / F(np.sqrt(a[1] * a[1] + a[2] * a[2]))
            c = np.array([0.0, a[2] * il, -a[1] * il], dtype=F)
        out[i, 0:3] = np.cross(c, a)
        out[i, 3:6] = c
    return out


def cube_mesh(to_world):
    """The `cube` shape: 24 vertices / 12 triangles with per-face normals and [0,1]^2 uvs, in the vertex / face /
    triangle order of the reference's tables (src/shapes/cube.cpp:22-28, 75-104) so that primitive indices and the
    per-face uv parameterisation agree: faces -y, +y, +x, +z, -x, -z; every face is the quad
    v0, v0+2a, v0+2a+2b, v0+2b with uvs (0,1), (1,1), (1,0), (0,0) and triangles (0,1,2), (3,0,2)."""
    faces = [  # (normal, first corner v0, edge direction a, edge direction b)
        ((0, -1, 0), (1, -1, -1), (0, 0, 1), (-1, 0, 0)),
        ((0, 1, 0), (1, 1, -1), (-1, 0, 0), (0, 0, 1)),
        ((1, 0, 0), (1, -1, -1), (0, 1, 0), (0, 0, 1)),
        ((0, 0, 1), (1, -1, 1), (0, 1, 0), (-1, 0, 0)),
        ((-1, 0, 0), (-1, -1, 1), (0, 1, 0), (0, 0, -1)),
        ((0, 0, -1), (1, 1, -1), (0, -1, 0), (-1, 0, 0)),
    ]
    pos, nrm, uv, idx = [], [], [], []
    m = np.asarray(to_world, dtype=np.float64)
    for n, v0, ea, eb in faces:
        n, v0, ea, eb = (np.array(x, dtype=np.float64) for x in (n, v0, ea, eb))
        base = len(pos)
        nn = xform_normal(m, n)
        nn = nn / np.linalg.norm(nn)
        for corner, tc in ((v0, (0, 1)), (v0 + 2 * ea, (1, 1)), (v0 + 2 * ea + 2 * eb, (1, 0)), (v0 + 2 * eb, (0, 0))):
            pos.append(xform_point(m, corner))
            nrm.append(nn)
            uv.append(tc)
        idx += [(base, base + 1, base + 2), (base + 3, base, base + 2)]
    return (np.array(pos, dtype=F), np.array(nrm, dtype=F), np.array(uv, dtype=F), np.array(idx, dtype=np.uint32))


def vertex_normals(positions, indices):
    """Area/angle-weighted smooth vertex normals for generated meshes (TriMesh::computeNormals
    semantics: angle-weighted face normals, src/librender/trimesh.cpp:570-680)."""
    p = np.asarray(positions, dtype=np.float64)
    idx = np.asarray(indices, dtype=np.int64).reshape(-1, 3)
    n = np.zeros_like(p)
    for k in range(3):
        a, b, c = p[idx[:, k]], p[idx[:, (k + 1) % 3]], p[idx[:, (k + 2) % 3]]
        e1, e2 = b - a, c - a
        fn = np.cross(e1, e2)
        ln = np.linalg.norm(fn, axis=1, keepdims=True)
        fn = np.divide(fn, ln, out=np.zeros_like(fn), where=ln > 0)
        cosang = np.sum(e1 * e2, axis=1) / np.maximum(np.linalg.norm(e1, axis=1) * np.linalg.norm(e2, axis=1), 1e-30)
        ang = np.arccos(np.clip(cosang, -1, 1))[:, None]
        np.add.at(n, idx[:, k], fn * ang)
    ln = np.linalg.norm(n, axis=1, keepdims=True)
    return np.divide(n, ln, out=np.zeros_like(n), where=ln > 0).astype(F)
