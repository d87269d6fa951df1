"""TEST INFRASTRUCTURE — ctypes views of the parity oracles.  NOT part of the product.

Two checkers live under oracle/:
  * libse_oracle.so      our plain-C restatement (oracle/se_oracle.c), built anywhere with gcc;
  * _ref/libref_*.so     the reference's OWN sources compiled in place from /root/reference
                         (oracle/Makefile `make ref`); built in the dev container, travels prebuilt.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF_DIR = os.path.join(HERE, "_ref")
REFERENCE_ROOT = os.environ.get("SE_REFERENCE_ROOT", "/root/reference")

_vp = C.c_void_p


def _p(a):
    return None if a is None else a.ctypes.data_as(_vp)


def build_oracle(force=False):
    so = os.path.join(HERE, "libse_oracle.so")
    src = os.path.join(HERE, "se_oracle.c")
    if force or not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
        subprocess.check_call(["make", "-s", "-C", HERE, "libse_oracle.so"])
    return so


def build_ref():
    """Compile oracle/_ref from the reference sources where they lie (only where they exist)."""
    if os.path.isdir(REFERENCE_ROOT):
        subprocess.check_call(["make", "-s", "-C", HERE, "ref", "REF=" + REFERENCE_ROOT])
        return True
    return False


def have_ref(name="libref_etsi.so"):
    return os.path.exists(os.path.join(REF_DIR, name))


def _features_ex(fn, x, mode16k):
    x = np.ascontiguousarray(x, np.int16)
    hop = 160 if mode16k else 80
    feat = np.zeros((len(x) // hop + 6, 15), np.float32)
    fn.restype = C.c_long
    fn.argtypes = [_vp, C.c_long, _vp, _vp, C.c_int]
    rows = fn(_p(x), len(x), _p(feat), None, (1 if mode16k else 0) | 2)
    return feat[:rows].copy()


class InfoStruct(C.Structure):
    _fields_ = [("produced", C.c_int), ("stage0_ran", C.c_int), ("stage1_ran", C.c_int),
                ("flag_vad", C.c_int), ("hang_over", C.c_int), ("nb_speech", C.c_int),
                ("nb_frame0", C.c_int), ("nb_frame1", C.c_int),
                ("sp_var", C.c_int), ("sp_spec", C.c_int), ("sp_mel", C.c_int), ("sp_vadns", C.c_int),
                ("mean_en", C.c_float), ("alfa_gf", C.c_float), ("low_snr", C.c_float), ("den_en", C.c_float),
                ("gains", C.c_float * 25)]


INFO_DTYPE = np.dtype([("produced", "i4"), ("stage0_ran", "i4"), ("stage1_ran", "i4"), ("flag_vad", "i4"),
                       ("hang_over", "i4"), ("nb_speech", "i4"), ("nb_frame0", "i4"), ("nb_frame1", "i4"),
                       ("sp_var", "i4"), ("sp_spec", "i4"), ("sp_mel", "i4"), ("sp_vadns", "i4"),
                       ("mean_en", "f4"), ("alfa_gf", "f4"), ("low_snr", "f4"), ("den_en", "f4"),
                       ("gains", "f4", (25,))])
assert INFO_DTYPE.itemsize == C.sizeof(InfoStruct)


class Oracle:
    """Plain-C restatement (oracle/se_oracle.c)."""

    def __init__(self, fast=False):
        if fast:
            so = os.path.join(HERE, "libse_oracle_fast.so")
            if not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(os.path.join(HERE, "se_oracle.c")):
                subprocess.check_call(["make", "-s", "-C", HERE, "libse_oracle_fast.so"])
        else:
            so = build_oracle()
        L = self.L = C.CDLL(so)
        L.so_ns_create.restype = _vp
        L.so_ns_create.argtypes = [C.c_int]
        L.so_ns_destroy.argtypes = [_vp]
        L.so_ns_block.argtypes = [_vp, _vp, _vp, _vp]
        L.so_ns_spectrum.restype = C.POINTER(C.c_float)
        L.so_ns_spectrum.argtypes = [_vp, C.c_int]
        L.so_ns_tables.argtypes = [_vp] * 7
        L.so_etsi_denoise.restype = C.c_long
        L.so_etsi_denoise.argtypes = [_vp, C.c_long, _vp, _vp, _vp, _vp, _vp]
        L.so_p16_func_wiener.argtypes = [_vp, _vp, C.c_int] + [_vp] * 7
        L.so_gammatone.argtypes = [_vp, _vp, C.c_float, C.c_float, C.c_float, C.c_long]
        L.so_haircell.argtypes = [_vp, _vp, C.c_long]
        L.so_subband.argtypes = [_vp, C.c_long, _vp]
        L.so_addnoise.argtypes = [_vp, _vp, _vp, C.c_long, C.c_int]
        L.so_resynth_num_frames.argtypes = [C.c_long, C.c_int]
        L.so_resynth.argtypes = [_vp, C.c_long, _vp, C.c_int, C.c_int, _vp, _vp]
        L.so_ibm_num_frames.restype = C.c_long
        L.so_ibm_num_frames.argtypes = [C.c_long]
        L.so_single_ibm.argtypes = [_vp, _vp, C.c_long, _vp, _vp]
        L.so_rfft.argtypes = [_vp, C.c_int, C.c_int]
        L.so_rfft_cxx.argtypes = [_vp, C.c_int, C.c_int]
        L.so_wave_proc.argtypes = [_vp]
        L.so_etsi_features.restype = C.c_long
        L.so_etsi_features.argtypes = [_vp, C.c_long, _vp, _vp]
        L.so_etsi_features_16k.restype = C.c_long
        L.so_etsi_features_16k.argtypes = [_vp, C.c_long, _vp, _vp]
        L.so_etsi_features_ex.restype = C.c_long
        L.so_etsi_features_ex.argtypes = [_vp, C.c_long, _vp, _vp, C.c_int]
        L.so_qmf_split.argtypes = [_vp, C.c_long, _vp, _vp]
        L.so_qmf_taps.argtypes = [_vp, _vp]
        L.so_ns_attach_hp16k.argtypes = [_vp]
        L.so_hp16k_tables.argtypes = [_vp, _vp, _vp, _vp]
        L.so_cc_dct26.argtypes = [_vp]

    # ---- primitives
    def rfft(self, x, m, cxx=False):
        y = np.ascontiguousarray(x, np.float32).copy()
        (self.L.so_rfft_cxx if cxx else self.L.so_rfft)(_p(y), len(y), m)
        return y

    # ---- P8
    def etsi_denoise(self, x, with_ceps=False, trace=False):
        x = np.ascontiguousarray(x, np.int16)
        n = len(x)
        den = np.zeros(n, np.int16)
        denf = np.zeros(n, np.float32)
        ncalls = n // 80
        ceps = np.zeros((max(ncalls, 1), 14), np.float32) if with_ceps else None
        nceps = C.c_long(0)
        info = np.zeros(max(ncalls, 1), INFO_DTYPE) if trace else None
        calls = self.L.so_etsi_denoise(_p(x), n, _p(den), _p(denf), _p(ceps), C.byref(nceps), _p(info))
        out = {"den_i16": den, "den_f32": denf, "calls": calls}
        if with_ceps:
            out["ceps"] = ceps[:nceps.value].copy()
        if trace:
            out["info"] = info[:calls]
        return out

    def ns_spectra(self, profile, x):
        """Block-by-block SPECTRUM state (so_ns_spectrum planes 0..7) behind the caller's zero-frame gate:
        profile 8 = DoAdvProcess (leading all-zero frames only, ParmInterface.c:244-250), 16 = func_Wiener (anywhere,
        NoiseSup.cpp:1159-1170).  Returns spectra[block][8][spec], ran[block] (bit 1 = stage 1 ran, bit 0 = stage 2
        ran, bit 2 = skipped by the gate) -- the layout of se_ns_spectra_batch."""
        shift, spec = (80, 65) if profile == 8 else (160, 129)
        xf = np.ascontiguousarray(x, np.float32)
        nb = len(xf) // shift
        h = self.L.so_ns_create(profile)
        sp = np.zeros((nb, 8, spec), np.float32)
        ran = np.zeros(nb, np.int32)
        out = np.zeros(shift, np.float32)
        info = np.zeros(1, INFO_DTYPE)
        onset = False
        for b in range(nb):
            blk = np.ascontiguousarray(xf[b * shift:(b + 1) * shift])
            nonzero = int(np.float32(np.sum(blk.astype(np.float64) ** 2))) != 0
            if profile == 8:
                if not (nonzero or onset):
                    ran[b] = 4
                    continue
                onset = True
            elif not nonzero:
                ran[b] = 4
                continue
            self.L.so_ns_block(h, _p(blk), _p(out), _p(info))
            ran[b] = (1 if info[0]["stage1_ran"] else 0) | (2 if info[0]["stage0_ran"] else 0)
            for st, bit in ((0, 2), (1, 1)):
                if ran[b] & bit:
                    for w in (3 * st, 3 * st + 1, 3 * st + 2, 6 + st):
                        sp[b, w] = np.ctypeslib.as_array(self.L.so_ns_spectrum(h, w), (spec,))
        self.L.so_ns_destroy(h)
        return sp, ran

    def comp_ceps(self, frame201):
        f = np.ascontiguousarray(frame201, np.float32)
        o = np.zeros(14, np.float32)
        self.L.so_comp_ceps(_p(f), _p(o))
        return o

    def wave_proc(self, frame200):
        f = np.ascontiguousarray(frame200, np.float32).copy()
        self.L.so_wave_proc(_p(f))
        return f

    def etsi_features(self, x):
        """NoiseSup -> WaveProc -> CompCeps -> PostProc -> VAD: (feat[rows][15], den_i16)."""
        x = np.ascontiguousarray(x, np.int16)
        feat = np.zeros((max(len(x) // 80, 1), 15), np.float32)
        den = np.zeros(len(x), np.int16)
        rows = self.L.so_etsi_features(_p(x), len(x), _p(feat), _p(den))
        return feat[:rows].copy(), den

    def etsi_features_flush(self, x, mode16k=False):
        return _features_ex(self.L.so_etsi_features_ex, x, mode16k)

    def etsi_features_16k(self, x):
        """16 kHz mode (QMF + HP-band coding): (feat[rows][15], low-band den_i16[(n//160)*80])."""
        x = np.ascontiguousarray(x, np.int16)
        feat = np.zeros((max(len(x) // 160, 1), 15), np.float32)
        den = np.zeros((len(x) // 160) * 80, np.int16)
        rows = self.L.so_etsi_features_16k(_p(x), len(x), _p(feat), _p(den))
        return feat[:rows].copy(), den

    def qmf_split(self, x):
        x = np.ascontiguousarray(x, np.int16)
        m = (len(x) // 160) * 80
        lp = np.zeros(m, np.float32); hp = np.zeros(m, np.float32)
        self.L.so_qmf_split(_p(x), len(x), _p(lp), _p(hp))
        return lp, hp

    def hp16k_tables(self):
        dp = np.zeros(118, np.float32); hp = np.zeros(118, np.float32)
        self.L.so_qmf_taps(_p(dp), _p(hp))
        h = self.L.so_ns_create(8)
        self.L.so_ns_attach_hp16k(h)
        st = np.zeros(5, np.int32); ln = np.zeros(5, np.int32); fb = np.zeros((5, 65), np.float32)
        self.L.so_hp16k_tables(h, _p(st), _p(ln), _p(fb))
        self.L.so_ns_destroy(h)
        dct26 = np.zeros((12, 26), np.float32)
        self.L.so_cc_dct26(_p(dct26))
        return dict(dp=dp, hp=hp, start=st, len=ln, fb=fb, dct26=dct26)

    def cc_tables(self):
        ham = np.zeros(100, np.float32); dct = np.zeros((12, 23), np.float32)
        st = np.zeros(23, np.int32); ln = np.zeros(23, np.int32); tri = np.zeros((23, 129), np.float32)
        self.L.so_cc_tables(_p(ham), _p(dct), _p(st), _p(ln), _p(tri))
        return dict(hamming=ham, dct=dct, start=st, len=ln, tri=tri)

    def ns_tables(self, profile):
        h = self.L.so_ns_create(profile)
        fl, w = (200, 65) if profile == 8 else (480, 128)
        sw = np.zeros(fl, np.float32); ir = np.zeros(17, np.float32); bank = np.zeros((25, w), np.float32)
        st = np.zeros(25, np.int32); ln = np.zeros(25, np.int32); idct = np.zeros((25, 25), np.float32)
        self.L.so_ns_tables(h, _p(sw), _p(ir), _p(bank), _p(st), _p(ln), _p(idct))
        self.L.so_ns_destroy(h)
        return dict(sig_window=sw, ir_window=ir, bank=bank, bank_start=st, bank_len=ln, idct=idct)

    # ---- P16
    def p16_stream(self):
        return _OracleP16(self)

    # ---- G64
    def channel_params(self):
        cf = np.zeros(64, np.float32); bw = np.zeros(64, np.float32); mid = np.zeros(64, np.float32)
        self.L.so_channel_params(_p(cf), _p(bw), _p(mid))
        return cf, bw, mid

    def gammatone(self, x, cf, bw, mid):
        x = np.ascontiguousarray(x, np.float32); y = np.zeros_like(x)
        self.L.so_gammatone(_p(x), _p(y), float(cf), float(bw), float(mid), len(x))
        return y

    def haircell(self, x):
        x = np.ascontiguousarray(x, np.float32); y = np.zeros_like(x)
        self.L.so_haircell(_p(x), _p(y), len(x))
        return y

    def subband(self, wav):
        wav = np.ascontiguousarray(wav, np.int16)
        out = np.zeros((64, len(wav)), np.int16)
        self.L.so_subband(_p(wav), len(wav), _p(out))
        return out

    def addnoise(self, pure, noise, db):
        pure = np.ascontiguousarray(pure, np.int16); noise = np.ascontiguousarray(noise, np.int16).copy()
        noisy = np.zeros_like(pure)
        self.L.so_addnoise(_p(pure), _p(noise), _p(noisy), len(pure), int(db))
        return noise, noisy

    def resynth_num_frames(self, n, mode):
        return self.L.so_resynth_num_frames(n, mode)

    def resynth(self, wav, mask, mode, want_f32=False):
        wav = np.ascontiguousarray(wav, np.int16)
        m = np.ascontiguousarray(mask, np.float32).copy()
        out = np.zeros(len(wav), np.int16)
        of = np.zeros(len(wav), np.float32) if want_f32 else None
        self.L.so_resynth(_p(wav), len(wav), _p(m), m.shape[0], mode, _p(out), _p(of))
        return (out, of, m) if want_f32 else (out, m)

    def single_ibm(self, pure64, noise64):
        pure64 = np.ascontiguousarray(pure64, np.int16); noise64 = np.ascontiguousarray(noise64, np.int16)
        n = pure64.shape[1]
        fr = self.L.so_ibm_num_frames(n)
        irm = np.zeros((fr, 64), np.float32); ibm = np.zeros((fr, 64), np.uint8)
        self.L.so_single_ibm(_p(pure64), _p(noise64), n, _p(irm), _p(ibm))
        return irm, ibm


class _OracleP16:
    def __init__(self, o):
        self.o = o
        self.h = o.L.so_ns_create(16)

    def close(self):
        if self.h:
            self.o.L.so_ns_destroy(self.h)
            self.h = None

    def func_wiener(self, x):
        x = np.ascontiguousarray(x, np.float32)
        n = len(x); nb = n // 160
        out = np.zeros(n, np.float32)
        fl = [np.zeros(max(nb, 1), np.int32) for _ in range(5)]
        gains = np.zeros((max(nb, 1), 25), np.float32)
        rows = self.o.L.so_p16_func_wiener(self.h, _p(x), n, _p(out), *[_p(f) for f in fl], _p(gains))
        return dict(out=out, sp_var=fl[0][:nb], sp_spec=fl[1][:nb], sp_mel=fl[2][:nb], sp_vadns=fl[3][:nb],
                    frame_counter=fl[4][:nb], gains=gains[:rows].copy())

    def spectrum(self, which, n=129):
        ptr = self.o.L.so_ns_spectrum(self.h, which)
        return np.ctypeslib.as_array(ptr, (n,)).copy()


# ------------------------------------------------------------------------------------------
# oracle/_ref : the reference's own code
# ------------------------------------------------------------------------------------------
class TraceStruct(C.Structure):
    _fields_ = [(k, _vp) for k in ("ret", "flag_vad", "hang_over", "nb_speech", "nb_frame0", "nb_frame1",
                                   "sp_var", "sp_spec", "sp_mel", "sp_vadns", "mean_en", "alfa_gf", "low_snr",
                                   "den_en", "nsig1", "noise1", "den1", "nsig2", "noise2", "den2", "out_f32")]


class RefEtsi:
    def __init__(self, fast=False):
        self.L = L = C.CDLL(os.path.join(REF_DIR, "libref_etsi_fast.so" if fast else "libref_etsi.so"))
        L.ref_etsi_denoise.argtypes = [_vp, _vp, C.c_long, C.c_int]
        L.ref_etsi_trace.restype = C.c_long
        L.ref_etsi_trace.argtypes = [_vp, C.c_long, _vp, _vp, _vp, _vp]
        L.ref_rfft.argtypes = [_vp, C.c_int, C.c_int]
        L.ref_comp_ceps.argtypes = [_vp, _vp]
        L.ref_wave_proc.argtypes = [_vp]
        L.ref_etsi_features.restype = C.c_long
        L.ref_etsi_features.argtypes = [_vp, C.c_long, _vp, _vp]
        L.ref_etsi_features_16k.restype = C.c_long
        L.ref_etsi_features_16k.argtypes = [_vp, C.c_long, _vp, _vp]
        L.ref_qmf_split.argtypes = [_vp, C.c_long, _vp, _vp]
        L.ref_adv_process_as_shipped.argtypes = [_vp, C.c_long, _vp, _vp, _vp, _vp, _vp]

    def rfft(self, x, m):
        y = np.ascontiguousarray(x, np.float32).copy()
        self.L.ref_rfft(_p(y), len(y), m)
        return y

    def adv_process_as_shipped(self, x):
        """DoAdvProcess frame by frame + the FlushAdvProcess loop, as shipped (feature block commented out):
        out int16[calls*80] (0x7a7a where DenoiseBuffer was not written), feat[calls][15] (7.0 where FeatureBuffer
        was not written), wrote[calls], flush_rets, flush_feat[k][15]."""
        x = np.ascontiguousarray(x, np.int16)
        calls = len(x) // 80
        out = np.zeros(max(calls, 1) * 80, np.int16)
        feat = np.zeros((max(calls, 1), 15), np.float32)
        wrote = np.zeros(max(calls, 1), np.int32)
        fr = np.zeros(8, np.int32)
        ff = np.zeros((8, 15), np.float32)
        k = self.L.ref_adv_process_as_shipped(_p(x), len(x), _p(out), _p(feat), _p(wrote), _p(fr), _p(ff))
        return dict(out=out[:calls * 80], feat=feat[:calls], wrote=wrote[:calls], flush_rets=fr[:k], flush_feat=ff[:k])

    def etsi_features_16k(self, x):
        """16 kHz mode: (feat[rows][15], low-band den_i16[(n//160)*80])."""
        x = np.ascontiguousarray(x, np.int16)
        feat = np.zeros((max(len(x) // 160, 1), 15), np.float32)
        den = np.zeros((len(x) // 160) * 80, np.int16)
        rows = self.L.ref_etsi_features_16k(_p(x), len(x), _p(feat), _p(den))
        return feat[:rows].copy(), den

    def etsi_features_flush(self, x, mode16k=False):
        """The feature chain + the ETSI driver's flush loop (while FlushAdvProcess(...) write): feat[rows][15]."""
        return _features_ex(self.L.ref_etsi_features_ex, x, mode16k)

    def qmf_split(self, x):
        x = np.ascontiguousarray(x, np.int16)
        m = (len(x) // 160) * 80
        lp = np.zeros(m, np.float32); hp = np.zeros(m, np.float32)
        self.L.ref_qmf_split(_p(x), len(x), _p(lp), _p(hp))
        return lp, hp

    def wave_proc(self, frame200):
        f = np.ascontiguousarray(frame200, np.float32).copy()
        self.L.ref_wave_proc(_p(f))
        return f

    def etsi_features(self, x):
        x = np.ascontiguousarray(x, np.int16)
        feat = np.zeros((max(len(x) // 80, 1), 15), np.float32)
        den = np.zeros(len(x), np.int16)
        rows = self.L.ref_etsi_features(_p(x), len(x), _p(feat), _p(den))
        return feat[:rows].copy(), den

    def etsi_denoise(self, x):
        x = np.ascontiguousarray(x, np.int16).copy()
        out = np.zeros_like(x)
        rc = self.L.ref_etsi_denoise(_p(x), _p(out), len(x), 1)
        return out, rc

    def trace(self, x, with_ceps=True):
        x = np.ascontiguousarray(x, np.int16)
        n = len(x); nc = max(n // 80, 1)
        den = np.zeros(n, np.int16)
        ceps = np.zeros((nc, 14), np.float32) if with_ceps else None
        nceps = C.c_long(0)
        ints = {k: np.zeros(nc, np.int32) for k in ("ret", "flag_vad", "hang_over", "nb_speech", "nb_frame0",
                                                    "nb_frame1", "sp_var", "sp_spec", "sp_mel", "sp_vadns")}
        fl = {k: np.zeros(nc, np.float32) for k in ("mean_en", "alfa_gf", "low_snr", "den_en")}
        sp = {k: np.zeros((nc, 65), np.float32) for k in ("nsig1", "noise1", "den1", "nsig2", "noise2", "den2")}
        outf = np.zeros((nc, 80), np.float32)
        tr = TraceStruct()
        for k, v in list(ints.items()) + list(fl.items()) + list(sp.items()) + [("out_f32", outf)]:
            setattr(tr, k, v.ctypes.data)
        calls = self.L.ref_etsi_trace(_p(x), n, _p(den), _p(ceps), C.byref(nceps), C.byref(tr))
        res = {"den_i16": den, "calls": calls, "out_f32": outf[:calls]}
        if with_ceps:
            res["ceps"] = ceps[:nceps.value].copy()
        for d in (ints, fl, sp):
            for k, v in d.items():
                res[k] = v[:calls]
        return res

    def comp_ceps(self, frame201):
        f = np.ascontiguousarray(frame201, np.float32); o = np.zeros(14, np.float32)
        self.L.ref_comp_ceps(_p(f), _p(o))
        return o

    def tables(self):
        sw = np.zeros(200, np.float32); ir = np.zeros(17, np.float32)
        ms = np.zeros(25, np.int32); ml = np.zeros(25, np.int32); md = np.zeros((25, 65), np.float32)
        idct = np.zeros((25, 25), np.float32); ham = np.zeros(100, np.float32); dct = np.zeros((12, 23), np.float32)
        cs = np.zeros(23, np.int32); cl = np.zeros(23, np.int32); cd = np.zeros((23, 129), np.float32)
        self.L.ref_etsi_tables(*[_p(a) for a in (sw, ir, ms, ml, md, idct, ham, dct, cs, cl, cd)])
        return dict(sig_window=sw, ir_window=ir, bank_start=ms, bank_len=ml, bank=md, idct=idct,
                    hamming=ham, dct=dct, cc_start=cs, cc_len=cl, cc_tri=cd)


class RefP16:
    def __init__(self, fast=False):
        self.L = L = C.CDLL(os.path.join(REF_DIR, "libref_p16_fast.so" if fast else "libref_p16.so"))
        L.ref_p16_create.restype = _vp
        L.ref_p16_destroy.argtypes = [_vp]
        L.ref_p16_func_wiener.argtypes = [_vp, _vp, C.c_int] + [_vp] * 8 + [_vp, C.c_long]
        L.ref_p16_state.argtypes = [_vp] * 9
        L.ref_p16_tables.argtypes = [_vp] * 6
        L.ref_p16_rfft.argtypes = [_vp, C.c_int, C.c_int]
        self.h = L.ref_p16_create()

    def close(self):
        if self.h:
            self.L.ref_p16_destroy(self.h)
            self.h = None

    def rfft(self, x, m):
        y = np.ascontiguousarray(x, np.float32).copy()
        self.L.ref_p16_rfft(_p(y), len(y), m)
        return y

    def func_wiener(self, x, want_text=False):
        x = np.ascontiguousarray(x, np.float32).copy()
        n = len(x); nb = n // 160
        out = np.zeros(n, np.float32)
        fl = [np.zeros(max(nb, 1), np.int32) for _ in range(5)]
        gains = np.zeros((max(nb, 1), 25), np.float32)
        rows = C.c_int(0)
        cap = 25 * 12 * max(nb, 1) + 16
        text = C.create_string_buffer(cap) if want_text else None
        self.L.ref_p16_func_wiener(self.h, _p(x), n, _p(out), *[_p(f) for f in fl], _p(gains), C.byref(rows),
                                   text, cap if want_text else 0)
        r = dict(out=out, sp_var=fl[0][:nb], sp_spec=fl[1][:nb], sp_mel=fl[2][:nb], sp_vadns=fl[3][:nb],
                 frame_counter=fl[4][:nb], gains=gains[:rows.value].copy())
        if want_text:
            r["text"] = text.value.decode()
        return r

    def state(self):
        sp = [np.zeros(129, np.float32) for _ in range(6)]
        sc = np.zeros(8, np.float32); isc = np.zeros(8, np.int32)
        self.L.ref_p16_state(self.h, *[_p(a) for a in sp], _p(sc), _p(isc))
        return sp, sc, isc

    def tables(self):
        sw = np.zeros(480, np.float32); ir = np.zeros(17, np.float32); g = np.zeros((25, 128), np.float32)
        gs = np.zeros(25, np.int32); idct = np.zeros((25, 25), np.float32)
        self.L.ref_p16_tables(self.h, _p(sw), _p(ir), _p(g), _p(gs), _p(idct))
        return dict(sig_window=sw, ir_window=ir, bank=g, bank_start=gs, idct=idct)


class RefG64:
    """variant in {'subband','resynth_ibm','resynth_ori','resynth_1dnn'}"""

    def __init__(self, variant, fast=False):
        name = "libref_%s%s.so" % (variant, "_fast" if fast else "")
        self.L = L = C.CDLL(os.path.join(REF_DIR, name))
        self.variant = variant
        L.ref_gammatone.argtypes = [_vp, _vp, C.c_float, C.c_float, C.c_float, C.c_long]
        L.ref_haircell.argtypes = [_vp, _vp, C.c_long]
        L.ref_channel_params.argtypes = [_vp, _vp, _vp]
        if variant == "subband":
            L.ref_subband.argtypes = [_vp, C.c_long, _vp]
            L.ref_addnoise.argtypes = [_vp, _vp, _vp, C.c_long, C.c_int]
        else:
            L.ref_resynth.argtypes = [_vp, C.c_long, _vp, C.c_int, _vp]

    def channel_params(self):
        cf = np.zeros(64, np.float32); bw = np.zeros(64, np.float32); mid = np.zeros(64, np.float32)
        self.L.ref_channel_params(_p(cf), _p(bw), _p(mid))
        return cf, bw, mid

    def gammatone(self, x, cf, bw, mid):
        x = np.ascontiguousarray(x, np.float32).copy(); y = np.zeros_like(x)
        self.L.ref_gammatone(_p(x), _p(y), float(cf), float(bw), float(mid), len(x))
        return y

    def haircell(self, x):
        x = np.ascontiguousarray(x, np.float32).copy(); y = np.zeros_like(x)
        self.L.ref_haircell(_p(x), _p(y), len(x))
        return y

    def subband(self, wav):
        wav = np.ascontiguousarray(wav, np.int16).copy()
        out = np.zeros((64, len(wav)), np.int16)
        rc = self.L.ref_subband(_p(wav), len(wav), _p(out))
        assert rc == 0
        return out

    def addnoise(self, pure, noise, db):
        pure = np.ascontiguousarray(pure, np.int16).copy(); noise = np.ascontiguousarray(noise, np.int16).copy()
        noisy = np.zeros_like(pure)
        self.L.ref_addnoise(_p(pure), _p(noise), _p(noisy), len(pure), int(db))
        return noise, noisy

    def resynth(self, wav, mask):
        wav = np.ascontiguousarray(wav, np.int16).copy()
        m = np.ascontiguousarray(mask, np.float32).copy()
        out = np.zeros(len(wav), np.int16)
        rc = self.L.ref_resynth(_p(wav), len(wav), _p(m), m.shape[0], _p(out))
        assert rc == 0
        return out, m


# ------------------------------------------------------------------------------------------
# oracle/ref_bench.c : the reference arm's timing loop (bench.py --impl reference / cpu_baseline)
# ------------------------------------------------------------------------------------------
class RefBench:
    """Times the reference's own CPU code (oracle/_ref/*_fast.so; the plain-C port where that is absent) over a batch
    of in-memory utterances on N forked worker processes -- no Python inside the timed region (oracle/ref_bench.c).
    workload: 'p8' (etsi_denoise + CompCeps step), 'p16' (func_Wiener), 'resynth' (resyth_64sub_IBM)."""
    LIBS = {"p8": "libref_etsi_fast.so", "p16": "libref_p16_fast.so", "resynth": "libref_resynth_ibm_fast.so"}
    IDS = {"p8": 0, "p16": 1, "resynth": 2}

    def __init__(self, workload):
        so = os.path.join(HERE, "libse_refbench.so")
        src = os.path.join(HERE, "ref_bench.c")
        if not os.path.exists(so) or (os.path.exists(src) and os.path.getmtime(so) < os.path.getmtime(src)):
            subprocess.check_call(["make", "-s", "-C", HERE, "libse_refbench.so"])
        self.L = C.CDLL(so)
        self.L.ref_bench_open.argtypes = [C.c_char_p]
        self.L.ref_bench_run.argtypes = [C.c_int, C.c_int, _vp, _vp, C.c_int, _vp, _vp, C.c_int,
                                         C.POINTER(C.c_double), C.POINTER(C.c_double)]
        self.workload = workload
        if have_ref(self.LIBS[workload]):
            self.kind, path = "reference", os.path.join(REF_DIR, self.LIBS[workload])
        else:
            self.kind = "port"
            path = os.path.join(HERE, "libse_oracle_fast.so")
            if not os.path.exists(path):
                subprocess.check_call(["make", "-s", "-C", HERE, "libse_oracle_fast.so"])
        self.lib_path = path
        if self.L.ref_bench_open(path.encode()) != 0:
            raise RuntimeError("cannot load " + path)

    def run(self, samples, offsets, workers, mask=None, frame_offsets=None):
        """-> (seconds, checksum) for utterances 0..len(offsets)-2."""
        samples = np.ascontiguousarray(samples, np.int16)
        offsets = np.ascontiguousarray(offsets, np.int64)
        if mask is not None:
            mask = np.ascontiguousarray(mask, np.float32)
            frame_offsets = np.ascontiguousarray(frame_offsets, np.int64)
        sec, cs = C.c_double(0), C.c_double(0)
        rc = self.L.ref_bench_run(self.IDS[self.workload], 1 if self.kind == "port" else 0, _p(samples), _p(offsets),
                                  len(offsets) - 1, _p(mask), _p(frame_offsets), int(workers), C.byref(sec), C.byref(cs))
        if rc != 0:
            raise RuntimeError("ref_bench_run failed (%d)" % rc)
        return sec.value, cs.value
