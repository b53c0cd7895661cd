"""ctypes front-end of oracle/liboracle.so (oracle.c) -- TEST INFRASTRUCTURE ONLY.

`Oracle` mirrors the observable semantics of the reference's `tomoengine` class
(/root/reference/tomofusion/gpu/utils/tomoengine.cpp) on the CPU so parity tests can run the same
call sequence against both.  It is never imported by the product package.
"""
import ctypes as C
import os
import subprocess
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

f32p = np.ctypeslib.ndpointer(dtype=np.float32, flags="C_CONTIGUOUS")
i32p = np.ctypeslib.ndpointer(dtype=np.int32, flags="C_CONTIGUOUS")


class AngleConst(C.Structure):
    _fields_ = [("angle", C.c_float), ("detsize", C.c_float), ("offset", C.c_float),
                ("sc", C.c_float), ("ss", C.c_float), ("off", C.c_float), ("scale", C.c_float),
                ("rayX", C.c_float), ("rayY", C.c_float), ("vertical", C.c_int32)]


def build():
    """Compile oracle.c -> liboracle.so (gcc, seconds)."""
    subprocess.check_call(["make", "-s", "-C", _HERE, "liboracle.so"])


def lib():
    global _LIB
    if _LIB is None:
        path = os.path.join(_HERE, "liboracle.so")
        if not os.path.exists(path) or os.path.getmtime(path) < os.path.getmtime(os.path.join(_HERE, "oracle.c")):
            build()
        _LIB = C.CDLL(path)
        L = _LIB
        vp = C.c_void_p
        L.orc_geometry.argtypes = [C.c_int, C.c_int, f32p, vp]
        L.orc_fp.argtypes = [f32p, f32p, C.c_int, C.c_int, C.c_int, C.c_int, vp, vp, vp, C.c_float, C.c_int]
        L.orc_bp.argtypes = [f32p, f32p, C.c_int, C.c_int, C.c_int, C.c_int, vp, C.c_float, C.c_int]
        L.orc_bp_sart.argtypes = [f32p, f32p, C.c_int, C.c_int, C.c_int, vp, C.c_float, C.c_int]
        L.orc_sirt_weights.argtypes = [f32p, f32p, C.c_int, C.c_int, C.c_int, C.c_int, vp, vp, vp, C.c_int]
        L.orc_fp3d.argtypes = [f32p, f32p, C.c_int, C.c_int, C.c_int, vp, vp, vp, C.c_int]
        L.orc_bp3d.argtypes = [f32p, f32p, C.c_int, C.c_int, C.c_int, vp, C.c_int]
        L.orc_sirt3d.argtypes = [f32p, f32p, f32p, f32p, C.c_int, C.c_int, C.c_int, vp, vp, vp, C.c_int, C.c_int]
        L.orc_sart3d.argtypes = [f32p, f32p, f32p, C.c_int, C.c_int, C.c_int, vp, vp, vp, vp, C.c_int, C.c_float, C.c_int]
        L.orc_lipschitz.argtypes = [C.c_int, C.c_int, vp, vp, vp, C.c_int]
        L.orc_lipschitz.restype = C.c_float
        L.orc_kl3d.argtypes = [f32p, f32p, C.c_int, C.c_int, C.c_int, vp, vp, vp, C.c_float, C.c_float, C.c_int]
        L.orc_kl3d.restype = C.c_float
        L.orc_tv.argtypes = [f32p, C.c_int, C.c_int, C.c_int]
        L.orc_tv.restype = C.c_double
        L.orc_tv_gradient.argtypes = [f32p, f32p, C.c_int, C.c_int, C.c_int]
        L.orc_tv_gd.argtypes = [f32p, C.c_int, C.c_float, C.c_int, C.c_int, C.c_int]
        L.orc_tv_gd.restype = C.c_float
        L.orc_tv_fgp.argtypes = [f32p, C.c_int, C.c_float, C.c_int, C.c_int, C.c_int]
        L.orc_tv_fgp.restype = C.c_float
        L.orc_momentum.argtypes = [f32p, f32p, f32p, C.c_float, C.c_size_t]
        L.orc_soft_threshold.argtypes = [f32p, C.c_float, C.c_size_t]
        L.orc_positivity.argtypes = [f32p, C.c_size_t]
        for nm in ("orc_sqdist",):
            getattr(L, nm).argtypes = [f32p, f32p, C.c_size_t]
            getattr(L, nm).restype = C.c_double
        for nm in ("orc_sum", "orc_l1", "orc_sumsq"):
            getattr(L, nm).argtypes = [f32p, C.c_size_t]
            getattr(L, nm).restype = C.c_double
        L.orc_num_threads.restype = C.c_int
    return _LIB


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class Geometry:
    """Per-angle constants (orc_geometry).  cosv/sinv optionally override cosf/sinf with the
    device-evaluated __cosf/__sinf tables so kernel-vs-oracle comparisons are not dominated by the
    fast-intrinsic error."""

    def __init__(self, angles_rad, ndet, cosv=None, sinv=None):
        self.angles = np.ascontiguousarray(angles_rad, dtype=np.float32)
        self.na = int(self.angles.size)
        self.nd = int(ndet)
        self.buf = (AngleConst * self.na)()
        lib().orc_geometry(self.na, self.nd, self.angles, C.cast(self.buf, C.c_void_p))
        self.cosv = None if cosv is None else np.ascontiguousarray(cosv, dtype=np.float32)
        self.sinv = None if sinv is None else np.ascontiguousarray(sinv, dtype=np.float32)

    @property
    def p(self):
        return C.cast(self.buf, C.c_void_p)

    def table(self):
        """(na, 10) float64 view of the constants for inspection."""
        return np.array([[getattr(g, f) for f, _ in AngleConst._fields_] for g in self.buf], dtype=np.float64)


def fp2d(vol, geo, outscale=1.0, quant=1, sino=None):
    H, W = vol.shape
    if sino is None:
        sino = np.zeros((geo.na, geo.nd), np.float32)
    lib().orc_fp(np.ascontiguousarray(vol, np.float32), sino, W, H, geo.na, geo.nd, geo.p,
                 _ptr(geo.cosv), _ptr(geo.sinv), outscale, quant)
    return sino


def bp2d(sino, geo, shape, outscale=1.0, quant=1, vol=None):
    H, W = shape
    if vol is None:
        vol = np.zeros((H, W), np.float32)
    lib().orc_bp(vol, np.ascontiguousarray(sino, np.float32), W, H, geo.na, geo.nd, geo.p, outscale, quant)
    return vol


class Oracle:
    """CPU mirror of tomoengine's state machine (tomoengine.cpp:48-484), float32, [x][y][z]."""

    def __init__(self, nslice, nray, angles_rad, quant=1, cosv=None, sinv=None):
        self.nx, self.n = int(nslice), int(nray)
        self.quant = int(quant)
        self.geo = Geometry(angles_rad, nray, cosv, sinv)
        self.na = self.geo.na
        self.recon = np.zeros((self.nx, self.n, self.n), np.float32)
        self.b = np.zeros((self.nx, self.na * self.n), np.float32)
        self.g = np.zeros_like(self.b)
        self.momentum = False
        self.yk = self.recon_old = self.temp = self.original = None
        self.R = self.C = None
        self.order = None
        self.relax = 1.0
        self.L_A = self.L_Aml = None

    # -- data ------------------------------------------------------------------------------
    def set_tilt_series(self, b):
        self.b = np.ascontiguousarray(b, dtype=np.float32).reshape(self.nx, self.na * self.n).copy()

    def set_original_volume(self, m, s):
        if self.original is None:
            self.original = np.zeros_like(self.recon)
        self.original[s] = m

    def initialize_initial_volume(self):
        self.original = np.zeros_like(self.recon)

    def initialize_recon_copy(self):
        self.temp = np.zeros_like(self.recon)

    def get_recon(self, s):
        return self.recon[s].copy()

    def set_recon(self, m, s):
        self.recon[s] = m

    def restart_recon(self):
        self.recon[:] = 0
        if self.momentum:
            self.yk[:] = 0
            self.recon_old[:] = 0

    # -- operators -------------------------------------------------------------------------
    def _weights(self):
        if self.R is None:
            self.R = np.zeros((self.na, self.n), np.float32)
            self.C = np.zeros((self.n, self.n), np.float32)
            lib().orc_sirt_weights(self.R, self.C, self.n, self.n, self.na, self.n, self.geo.p,
                                   _ptr(self.geo.cosv), _ptr(self.geo.sinv), self.quant)

    def initialize_SIRT(self):
        self._weights()

    def initialize_SART(self, order="sequential"):
        self._weights()
        self.order = np.arange(self.na, dtype=np.int32)
        self.sart_order_name = order

    def initialize_FP(self):
        pass

    def initialize_BP(self):
        pass

    def lipschitz(self):
        return lib().orc_lipschitz(self.n, self.na, self.geo.p, _ptr(self.geo.cosv), _ptr(self.geo.sinv), self.quant)

    def initialize_fista(self):
        self.momentum = True
        self.yk = self.recon.copy()
        self.recon_old = self.recon.copy()
        self._weights()
        self.L_A = self.lipschitz()

    def get_lipschitz(self):
        return self.L_A

    def remove_momentum(self):
        self.momentum = False

    def initialize_poisson_ML(self):
        self.L_Aml = self.lipschitz()
        m = self.b.max()
        if m > 1:
            self.b = (self.b / m).astype(np.float32)

    def SIRT(self, niter):
        tgt = self.yk if self.momentum else self.recon
        lib().orc_sirt3d(tgt, self.b, self.R, self.C, self.nx, self.n, self.na, self.geo.p,
                         _ptr(self.geo.cosv), _ptr(self.geo.sinv), niter, self.quant)

    def SART(self, beta, niter):
        if beta != 1:
            self.relax = float(beta)      # sticky, tomoengine.cpp:172
        lib().orc_sart3d(self.recon, self.b, self.R, self.nx, self.n, self.na, self.geo.p,
                         _ptr(self.geo.cosv), _ptr(self.geo.sinv), _ptr(self.order), niter, self.relax, self.quant)

    def poisson_ML(self, lam):
        return lib().orc_kl3d(self.recon, self.b, self.nx, self.n, self.na, self.geo.p,
                              _ptr(self.geo.cosv), _ptr(self.geo.sinv), lam, self.L_Aml, self.quant)

    def forward_projection(self):
        lib().orc_fp3d(self.recon, self.g, self.nx, self.n, self.na, self.geo.p,
                       _ptr(self.geo.cosv), _ptr(self.geo.sinv), self.quant)

    def create_projections(self):
        lib().orc_fp3d(self.original, self.b, self.nx, self.n, self.na, self.geo.p,
                       _ptr(self.geo.cosv), _ptr(self.geo.sinv), self.quant)

    def back_projection_of(self, sino):
        out = np.zeros_like(self.recon)
        lib().orc_bp3d(out, np.ascontiguousarray(sino, np.float32), self.nx, self.n, self.na, self.geo.p, self.quant)
        return out

    def data_distance(self):
        self.forward_projection()
        return float(np.sqrt(lib().orc_sqdist(self.g, self.b, self.g.size)))

    def get_projections(self):
        return self.b.copy()

    def get_model_projections(self):
        return self.g.copy()

    # -- regularisers / elementwise -----------------------------------------------------------
    def tv(self):
        return float(lib().orc_tv(self.recon, self.nx, self.n, self.n))

    def original_tv(self):
        return float(lib().orc_tv(self.original, self.nx, self.n, self.n))

    def tv_gd(self, ng, dPOCS):
        return lib().orc_tv_gd(self.recon, ng, dPOCS, self.nx, self.n, self.n)

    def tv_fgp(self, ng, lam):
        return lib().orc_tv_fgp(self.recon, ng, lam, self.nx, self.n, self.n)

    def fista_momentum(self, beta):
        lib().orc_momentum(self.recon, self.yk, self.recon_old, beta, self.recon.size)

    def copy_recon(self):
        self.temp[:] = self.recon

    def matrix_2norm(self):
        return float(np.sqrt(lib().orc_sqdist(self.recon, self.temp, self.recon.size)))

    def rmse(self):
        return float(np.sqrt(lib().orc_sqdist(self.recon, self.original, self.recon.size) / self.recon.size))

    def l1_norm(self):
        return float(lib().orc_l1(self.recon, self.recon.size))

    def soft_threshold(self, lam):
        tgt = self.yk if self.momentum else self.recon
        lib().orc_soft_threshold(tgt.reshape(-1), lam, tgt.size)
        lib().orc_positivity(tgt.reshape(-1), tgt.size)

    # -- CGLS: A/cuda/2d/cgls.cu:112-176, one independent run per slice (tomoengine.cpp:214-229) ----
    def initialize_CGLS(self):
        pass

    def CGLS(self, niter):
        f32 = np.float32
        x = self.recon
        fp = lambda v: self._fp_of(v)
        r = (self.b - fp(x)).astype(f32)
        p = self.back_projection_of(r)
        gamma = np.sum(p.astype(np.float64) ** 2, axis=(1, 2)).astype(f32)
        for _ in range(niter):
            w = fp(p)
            ww = np.sum(w.astype(np.float64) ** 2, axis=1).astype(f32)
            alpha = (gamma / ww).astype(f32)
            x += alpha[:, None, None] * p
            r = (r - alpha[:, None] * w).astype(f32)
            z = self.back_projection_of(r)
            gnew = np.sum(z.astype(np.float64) ** 2, axis=(1, 2)).astype(f32)
            beta = ((f32(1) / gamma) * gnew).astype(f32)
            gamma = gnew
            p = (z + beta[:, None, None] * p).astype(f32)
        np.maximum(x, 0, out=x)

    def _fp_of(self, vol):
        out = np.zeros_like(self.b)
        lib().orc_fp3d(np.ascontiguousarray(vol, np.float32), out, self.nx, self.n, self.na, self.geo.p,
                       _ptr(self.geo.cosv), _ptr(self.geo.sinv), self.quant)
        return out

    # -- FBP/WBP: A/cuda/2d/fbp.cu:105-190, fft.cu:251-375, A/src/Filters.cpp:40-420 ---------------
    def initialize_FBP(self, name):
        self.fbp_filter = name

    def FBP(self, positivity):
        n, na = self.n, self.na
        P = 1
        while P < 2 * n:
            P *= 2
        sino = self.b.reshape(self.nx, na, n)
        filt = fbp_filter(self.fbp_filter, P)
        if filt is not None:
            pad = np.zeros((self.nx, na, P), np.float32)
            pad[:, :, :n] = sino
            spec = np.fft.rfft(pad.astype(np.float64), axis=2) * filt.astype(np.float64)
            sino = (np.fft.irfft(spec, n=P, axis=2)[:, :, :n]).astype(np.float32)     # irfft already divides by P
        scale = np.float32((np.pi / np.float32(2.0)) / np.float32(na))
        out = np.zeros_like(self.recon)
        lib().orc_bp3d(out, np.ascontiguousarray(sino.reshape(self.nx, na * n), np.float32), self.nx, self.n, self.na, self.geo.p, self.quant)
        self.recon[:] = out * scale
        if positivity:
            np.maximum(self.recon, 0, out=self.recon)


def fbp_filter(name, P):
    """astra::genFilter (A/src/Filters.cpp:40-420) with m_fD = 1, m_fParameter = -1 (Filters.h:79);
    float32 window arithmetic, the ramp spectrum by a double-precision FFT.  None for 'none'."""
    from scipy.special import j0
    f32 = np.float32
    name = name.lower()
    if name == "none":
        return None
    F = P // 2 + 1
    PI = f32(3.14159265358979323846264338328)
    h = np.zeros(P, np.float32)
    i = np.arange(P)
    j = np.where(2 * i > P, P - i, i).astype(np.float32)
    odd = (i & 1) == 1
    fpi = (PI * j[odd]).astype(f32)
    h[odd] = f32(-1) / (fpi * fpi)
    h[0] = 0.25
    filt = (f32(2) * np.real(np.fft.fft(h.astype(np.float64)))[:F].astype(f32)).astype(f32)
    k = np.arange(F)
    w = (PI * f32(2) * (k.astype(f32) / f32(P))).astype(f32)
    n = k.astype(f32)
    N1 = f32(F - 1)

    def cos4(a0, a1, a2, a3, a4=0.0):
        base = (PI * n / N1).astype(f32)
        return (f32(a0) - f32(a1) * np.cos(f32(2) * base) + f32(a2) * np.cos(f32(4) * base) - f32(a3) * np.cos(f32(6) * base)
                + f32(a4) * np.cos(f32(8) * base)).astype(f32)

    with np.errstate(all="ignore"):
        if name == "ram-lak":
            v = np.ones(F, f32)
        elif name == "shepp-logan":
            v = np.sin(w / f32(2)) / (w / f32(2))
        elif name == "cosine":
            v = np.cos(w / f32(2))
        elif name == "hamming":
            v = f32(0.54) + f32(0.46) * np.cos(w)
        elif name == "hann":
            v = (f32(1) + np.cos(w)) / f32(2)
        elif name == "tukey":
            half = f32(F) / f32(2); en = f32(0.5) * half; den = f32(0.5) * half
            v = np.where((half - en <= n) & (n <= half + en), f32(1), f32(0.5) * (f32(1) + np.cos(PI * (n - en) / den)))
        elif name == "lanczos":
            s = (PI * (f32(2) * n / N1 - f32(1))).astype(f32)
            v = np.where(np.abs(s) > f32(0.001), np.sin(s.astype(np.float64)) / s.astype(np.float64), 1.0)
        elif name == "triangular":
            v = f32(2) / N1 * (N1 / f32(2) - np.abs(n - N1 / f32(2)))
        elif name == "gaussian":
            q = (f32(F) - f32(1)) / f32(2); t = (n - q) / (f32(0.4) * q)
            v = np.exp(f32(-0.5) * t * t)
        elif name == "barlett-hann":
            v = f32(0.62) - f32(0.48) * np.abs(n / N1 - f32(0.5)) - f32(0.38) * np.cos(f32(2) * PI * n / N1)
        elif name == "blackman":
            al = f32(0.16)
            v = (f32(1) - al) / f32(2) - f32(0.5) * np.cos(f32(2) * PI * f32(0.5) * n / N1) + (al / f32(2)) * np.cos(f32(4) * PI * f32(0.5) * n / N1)
        elif name == "nuttall":
            v = cos4(0.355768, 0.487396, 0.144232, 0.012604)
        elif name == "blackman-harris":
            v = cos4(0.35875, 0.48829, 0.14128, 0.01168)
        elif name == "blackman-nuttall":
            v = cos4(0.3635819, 0.4891775, 0.1365995, 0.0106411)
        elif name == "flat-top":
            v = cos4(1.0, 1.93, 1.29, 0.388, 0.032)
        elif name == "kaiser":
            pa = PI * f32(3); sq = f32(2) * n / N1 - f32(1)
            v = j0((pa * np.sqrt((f32(1) - sq * sq).astype(np.float64)).astype(f32)).astype(np.float64)).astype(f32) / f32(j0(float(pa)))
        elif name == "parzen":
            q = n / N1
            v = np.where(q <= f32(0.5), f32(1) - f32(6) * q * q * (f32(1) - q), f32(2) * (f32(1) - q) ** 3)
        else:
            raise ValueError(f"unknown FBP filter {name}")
    v = np.asarray(v, np.float32); v[0] = 1.0
    filt = (filt * v).astype(f32)
    filt[w > PI] = 0
    return filt


class OracleMM:
    """CPU mirror of the reference's `multimodal` state machine
    (/root/reference/tomofusion/chemistry/utils/multimodal.cpp), float32, recon[e][x][y][z].
    Projector calls go through oracle.c; the Eigen vector arithmetic is restated with numpy float32
    operations in the same order (one rounding per operation, no contraction)."""

    def __init__(self, ns, nray, nel, haadf_angles, chem_angles, quant=1, trig_h=(None, None), trig_c=(None, None)):
        self.nx, self.n, self.nel = int(ns), int(nray), int(nel)
        self.oh = Oracle(ns, nray, haadf_angles, quant, *trig_h)      # HAADF geometry; oh.b = bh, oh.g = g
        self.oc = Oracle(ns, nray, chem_angles, quant, *trig_c)       # chemical geometry
        self.nh, self.nc = self.oh.na, self.oc.na
        self.X = np.zeros((self.nel, self.nx, self.n, self.n), np.float32)
        self.bc = np.zeros((self.nel, self.nx, self.nc * self.n), np.float32)
        self.W = None
        self.gamma = np.float32(1.0)
        self.eps = np.float32(1e-1)
        self.measureHaadf = self.measureChem = False
        self.L_Aps = None

    # multimodal.cpp:156,159
    def set_haadf_tilt_series(self, bh):
        self.oh.set_tilt_series(bh)

    def set_chem_tilt_series(self, bchem):
        b = np.ascontiguousarray(bchem, np.float32).reshape(self.nx, self.nel, self.nc * self.n)
        self.bc = np.ascontiguousarray(np.transpose(b, (1, 0, 2)))

    def set_gamma(self, g):
        self.gamma = np.float32(g)

    def set_measureHaadf(self, f):
        self.measureHaadf = bool(f)

    def set_measureChem(self, f):
        self.measureChem = bool(f)

    # multimodal.cpp:268-274 with the structure of fusion_helper.py:5-32: entry (p, p + Npix*e) = w
    def load_sigma(self, trip):
        t = np.asarray(trip, np.float32)
        npix = self.n * self.n
        self.W = np.zeros((self.nel, npix), np.float32)
        r, c = t[0].astype(np.int64), t[1].astype(np.int64)
        assert np.all(c % npix == r)
        self.W[c // npix, r] = t[2]
        self.W = self.W.reshape(self.nel, 1, self.n, self.n)

    def initialize_FP(self): pass
    def initialize_BP(self): pass

    def initialize_SIRT(self):
        self.oh.initialize_SIRT()

    def estimate_lipschitz(self):            # :259-265
        self.L_Aps = np.float32(self.oc.lipschitz())

    def restart_recon(self):
        self.X[:] = 0

    def get_recon(self, e, s):
        return self.X[e, s].copy()

    def _model(self):                        # sigma * x^gamma, row-major sparse product: one rounded product added at a time
        m = np.zeros((self.nx, self.n, self.n), np.float32)
        for e in range(self.nel):
            v = self.X[e] if self.gamma == 1 else np.power(self.X[e], self.gamma, dtype=np.float32)
            m = m + self.W[e] * v
        return m

    def _fp(self, o, vol):
        out = np.zeros((self.nx, o.na * self.n), np.float32)
        lib().orc_fp3d(np.ascontiguousarray(vol, np.float32), out, self.nx, self.n, o.na, o.geo.p, _ptr(o.geo.cosv), _ptr(o.geo.sinv), o.quant)
        return out

    def poisson_ml(self, lam):               # :277-304
        cost = 0.0
        for e in range(self.nel):
            x = np.ascontiguousarray(self.X[e])
            cost += lib().orc_kl3d(x, np.ascontiguousarray(self.bc[e]), self.nx, self.n, self.nc, self.oc.geo.p,
                                   _ptr(self.oc.geo.cosv), _ptr(self.oc.geo.sinv), lam, self.L_Aps, self.oc.quant)
            self.X[e] = x
        return float(cost) if self.measureChem else 0.0

    def rescale_tomograms(self, scale):      # :307-309
        self.X *= np.float32(scale)

    def rescale_projections(self):           # :312-328
        self.oh.g = self._fp(self.oh, self._model())
        b = self.oh.b.reshape(self.nx, self.nh, self.n)
        g = self.oh.g.reshape(self.nx, self.nh, self.n)
        for p in range(self.nh):
            b[:, p, :] = b[:, p, :] / b[:, p, :].max()
            b[:, p, :] = b[:, p, :] * g[:, p, :].max()

    def sirt_data_fusion(self, lamH, lamC, niter):   # data_fusion("SIRT"), :452-491; fuse :425-441
        lamH, lamC = np.float32(lamH), np.float32(lamC)
        m = self._model()
        self.oh.g = self._fp(self.oh, m)
        u = m.copy()
        self.oh._weights()
        lib().orc_sirt3d(u, self.oh.b, self.oh.R, self.oh.C, self.nx, self.n, self.nh, self.oh.geo.p,
                         _ptr(self.oh.geo.cosv), _ptr(self.oh.geo.sinv), niter, self.oh.quant)
        c = np.float32(lamC / self.L_Aps)
        d = u - m
        cost = 0.0
        for e in range(self.nel):
            x = self.X[e]
            Ax = self._fp(self.oc, x)
            b = self.bc[e]
            cost += float(np.sum((Ax - b * np.log(Ax + self.eps)).astype(np.float64)))
            ratio = (Ax - b) / (Ax + self.eps)
            updC = self.oc.back_projection_of(ratio)
            h = self.W[e] * d
            if self.gamma != 1:
                h = (self.gamma * np.power(x, self.gamma - np.float32(1), dtype=np.float32)) * h
            xn = x - (c * updC - lamH * h)
            self.X[e] = np.maximum(xn, np.float32(0))
        costH = float(np.sqrt(lib().orc_sqdist(self.oh.g, self.oh.b, self.oh.g.size))) if self.measureHaadf else 0.0
        return costH, (cost if self.measureChem else 0.0)

    def data_distance(self):                 # :207-218
        s = 0.0
        for e in range(self.nel):
            g = self._fp(self.oc, self.X[e])
            s += lib().orc_sqdist(g, np.ascontiguousarray(self.bc[e]), g.size)
        return float(np.sqrt(s))

    def tv_fgp_4D(self, ng, lam):            # chemistry regularizers/tv_fgp.cu:192-281 (element loop)
        tv = np.float32(0)
        for e in range(self.nel):
            x = np.ascontiguousarray(self.X[e])
            tv = np.float32(tv + np.float32(lib().orc_tv_fgp(x, ng, lam, self.nx, self.n, self.n)))
            self.X[e] = x
        return float(tv)

    def tv_gd(self, ng, dPOCS):
        tv = np.float32(0)
        for e in range(self.nel):
            x = np.ascontiguousarray(self.X[e])
            tv = np.float32(tv + np.float32(lib().orc_tv_gd(x, ng, dPOCS, self.nx, self.n, self.n)))
            self.X[e] = x
        return float(tv)

    def get_haadf_projections(self):
        return self.oh.b.copy()

    def get_model_projections(self):
        return self.oh.g.copy()

    def get_chem_projections(self):
        return np.ascontiguousarray(np.transpose(self.bc, (1, 0, 2)).reshape(self.nx, -1))
