"""ctypes bindings for the two CPU checkers (TEST INFRASTRUCTURE).

* oracle/liblora_oracle.so      -- plain-C restatement (prefix ``oracle_``)
* oracle/_ref/libref_lora_phy.so -- the reference's own sources (prefix ``ref_``)

Both export the same call shapes, so one binder serves both; ``CpuPhy`` hides the
prefix.  Nothing in the product package imports this module.
"""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_SO = os.path.join(ROOT, "oracle", "liblora_oracle.so")
REF_SO = os.path.join(ROOT, "oracle", "_ref", "libref_lora_phy.so")

u8p = C.POINTER(C.c_uint8)
u16p = C.POINTER(C.c_uint16)
f32p = C.POINTER(C.c_float)
i32p = C.POINTER(C.c_int)


def _p(a, t):
    return a.ctypes.data_as(t)


class ChainCfg(C.Structure):
    _fields_ = [("sf", C.c_uint), ("ppm", C.c_uint), ("rdd", C.c_uint), ("explicit_hdr", C.c_int), ("crc", C.c_int),
                ("whitening", C.c_int), ("error_check", C.c_int), ("crc_check", C.c_int), ("out_hdr", C.c_int),
                ("data_length", C.c_uint)]


class CpuPhy:
    def __init__(self, lib, prefix):
        self.lib, self.prefix = lib, prefix
        self.is_ref = prefix == "ref_"

    def fn(self, name, restype, argtypes):
        f = getattr(self.lib, self.prefix + name)
        f.restype, f.argtypes = restype, argtypes
        return f

    # ---- demod ---------------------------------------------------------------------
    def lora_demodulate(self, iq, sf, osr=1, window=0, with_scratch=1):
        iq = np.ascontiguousarray(iq, dtype=np.complex64)
        n = iq.size
        nsym = n // ((1 << sf) * osr)
        out = np.zeros(max(nsym, 1), dtype=np.uint16)
        sync = C.c_uint8(0)
        cfo, to = C.c_float(0), C.c_float(0)
        f = self.fn("lora_demodulate", C.c_long,
                    [C.c_uint, C.c_int, C.c_int, C.c_void_p, C.c_size_t, C.c_uint, u16p,
                     C.POINTER(C.c_uint8), f32p, f32p])
        r = f(sf, window, with_scratch, iq.ctypes.data, n, osr, _p(out, u16p),
              C.byref(sync), C.byref(cfo), C.byref(to))
        cnt = max(int(r), 0)
        return int(r), out[:cnt].copy(), sync.value, np.float32(cfo.value), np.float32(to.value)

    def detect(self, x, sf):
        x = np.ascontiguousarray(x, dtype=np.complex64)
        N = 1 << sf
        spec = np.zeros(N, dtype=np.complex64)
        p, pav, fi = C.c_float(), C.c_float(), C.c_float()
        if self.is_ref:
            f = self.fn("detect", C.c_long, [C.c_uint, C.c_void_p, C.c_void_p, f32p, f32p, f32p])
            idx = f(sf, x.ctypes.data, spec.ctypes.data, C.byref(p), C.byref(pav), C.byref(fi))
        else:
            f = self.fn("detect", C.c_size_t, [C.c_uint, C.c_void_p, C.c_void_p, f32p, f32p, f32p])
            idx = f(sf, x.ctypes.data, spec.ctypes.data, C.byref(p), C.byref(pav), C.byref(fi))
        return int(idx), spec, np.float32(p.value), np.float32(pav.value), np.float32(fi.value)

    def twiddles(self, sf):
        out = np.zeros(1 << sf, dtype=np.complex64)
        self.fn("twiddles", None, [C.c_uint, C.c_void_p])(sf, out.ctypes.data)
        return out

    # ---- modulator -----------------------------------------------------------------
    def lora_modulate(self, symbols, sf, osr=1, bw_hz=125000, amplitude=1.0, sync=0x12):
        symbols = np.ascontiguousarray(symbols, dtype=np.uint16)
        n = symbols.size
        out = np.zeros((n + 2) * (1 << sf) * osr, dtype=np.complex64)
        f = self.fn("lora_modulate", C.c_size_t,
                    [C.c_void_p, C.c_size_t, C.c_void_p, C.c_uint, C.c_uint, C.c_uint, C.c_float, C.c_uint8])
        r = f(symbols.ctypes.data if n else None, n, out.ctypes.data, sf, osr, bw_hz, amplitude, sync)
        assert r == out.size
        return out

    def gen_chirp(self, N, osr, NN, f0, down, ampl, phase, bw_scale):
        out = np.zeros(NN, dtype=np.complex64)
        ph = C.c_float(phase)
        f = self.fn("gen_chirp", C.c_int,
                    [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_float, C.c_int, C.c_float, f32p, C.c_float])
        f(out.ctypes.data, N, osr, NN, f0, int(down), ampl, C.byref(ph), bw_scale)
        return out, np.float32(ph.value)

    def downchirp(self, sf, bw_hz=125000):
        N = 1 << sf
        return self.gen_chirp(N, 1, N, 0.0, True, 1.0, 0.0, bw_hz / 125000.0)[0]

    # ---- shipped coder -------------------------------------------------------------
    def lora_encode(self, payload, sf=7):
        payload = np.ascontiguousarray(payload, dtype=np.uint8)
        out = np.zeros(2 * payload.size, dtype=np.uint16)
        if self.is_ref:
            f = self.fn("lora_encode", C.c_size_t, [C.c_void_p, C.c_size_t, C.c_void_p, C.c_uint])
            f(payload.ctypes.data, payload.size, out.ctypes.data, sf)
        else:
            f = self.fn("lora_encode", C.c_size_t, [C.c_void_p, C.c_size_t, C.c_void_p])
            f(payload.ctypes.data, payload.size, out.ctypes.data)
        return out

    def lora_decode(self, symbols):
        symbols = np.ascontiguousarray(symbols, dtype=np.uint16)
        out = np.zeros(symbols.size // 2 + 1, dtype=np.uint8)
        f = self.fn("lora_decode", C.c_long, [C.c_void_p, C.c_size_t, C.c_void_p])
        r = f(symbols.ctypes.data, symbols.size, out.ctypes.data)
        return int(r), out[:max(int(r), 0)].copy()

    # ---- codes ---------------------------------------------------------------------
    def data_checksum(self, data):
        data = np.ascontiguousarray(data, dtype=np.uint8)
        return self.fn("data_checksum", C.c_uint16, [C.c_void_p, C.c_int])(data.ctypes.data, data.size)

    def header_checksum(self, h):
        h = np.ascontiguousarray(h, dtype=np.uint8)
        return self.fn("header_checksum", C.c_uint8, [C.c_void_p])(h.ctypes.data)

    def whiten(self, kind, data, ofs=0, rdd=4):
        b = np.array(data, dtype=np.uint8, copy=True)
        if kind == "sx1232":
            self.fn("whiten_sx1232", None, [C.c_void_p, C.c_uint16])(b.ctypes.data, b.size)
        elif kind == "table":
            self.fn("whiten_table", None, [C.c_void_p, C.c_uint16, C.c_int, C.c_int])(b.ctypes.data, b.size, ofs, rdd)
        else:
            self.fn("whiten_lfsr", None, [C.c_void_p, C.c_uint16, C.c_int, C.c_size_t])(b.ctypes.data, b.size, ofs, rdd)
        return b

    def u8fn(self, name, v):
        return self.fn(name, C.c_uint8, [C.c_uint8])(v)

    def u8fn_err(self, name, v, two=False):
        e, b = C.c_int(0), C.c_int(0)
        if two:
            r = self.fn(name, C.c_uint8, [C.c_uint8, i32p, i32p])(v, C.byref(e), C.byref(b))
            return r, e.value, b.value
        r = self.fn(name, C.c_uint8, [C.c_uint8, i32p])(v, C.byref(e))
        return r, e.value

    def gray(self, name, v):
        return self.fn(name, C.c_uint16, [C.c_uint16])(v)

    def interleave(self, cw, ppm, rdd):
        cw = np.ascontiguousarray(cw, dtype=np.uint8)
        out = np.zeros((cw.size // ppm) * (4 + rdd), dtype=np.uint16)
        self.fn("interleave", None, [C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t, C.c_size_t])(
            cw.ctypes.data, cw.size, out.ctypes.data, ppm, rdd)
        return out

    def deinterleave(self, sym, ppm, rdd):
        sym = np.ascontiguousarray(sym, dtype=np.uint16)
        out = np.zeros((sym.size // (4 + rdd)) * ppm, dtype=np.uint8)
        self.fn("deinterleave", None, [C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t, C.c_size_t])(
            sym.ctypes.data, sym.size, out.ctypes.data, ppm, rdd)
        return out

    # ---- high-level API (oracle restatement only; the reference side is RefWorkspace) -----
    def hl_demodulate(self, iq, sf, bw_hz=125000, osr=1, window=0, cap=None):
        assert not self.is_ref
        iq = np.ascontiguousarray(iq, dtype=np.complex64)
        cap = iq.size if cap is None else cap
        out = np.zeros(max(cap, 1), dtype=np.uint16)
        sync = C.c_uint8(0)
        cfo, to = C.c_float(0), C.c_float(0)
        f = self.fn("demodulate", C.c_long,
                    [C.c_uint, C.c_uint, C.c_uint, C.c_int, C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t,
                     C.POINTER(C.c_uint8), f32p, f32p])
        r = f(sf, bw_hz, osr, window, iq.ctypes.data, iq.size, out.ctypes.data, cap, C.byref(sync),
              C.byref(cfo), C.byref(to))
        return int(r), out[:max(int(r), 0)].copy(), sync.value, np.float32(cfo.value), np.float32(to.value)

    def hl_estimate_offsets(self, iq, sf, osr=1, window=0):
        assert not self.is_ref
        iq = np.ascontiguousarray(iq, dtype=np.complex64)
        cfo, to = C.c_float(0), C.c_float(0)
        f = self.fn("estimate_offsets", None, [C.c_uint, C.c_uint, C.c_int, C.c_void_p, C.c_size_t, f32p, f32p])
        f(sf, osr, window, iq.ctypes.data, iq.size, C.byref(cfo), C.byref(to))
        return np.float32(cfo.value), np.float32(to.value)

    def hl_compensate_offsets(self, iq, sf, osr, cfo, to):
        assert not self.is_ref
        iq = np.array(iq, dtype=np.complex64, copy=True)
        f = self.fn("compensate_offsets", None, [C.c_uint, C.c_uint, C.c_float, C.c_float, C.c_void_p, C.c_size_t])
        f(sf, osr, cfo, to, iq.ctypes.data, iq.size)
        return iq

    def hl_decode(self, symbols, cap=None):
        assert not self.is_ref
        symbols = np.ascontiguousarray(symbols, dtype=np.uint16)
        cap = symbols.size if cap is None else cap
        out = np.zeros(symbols.size // 2 + 8, dtype=np.uint8)
        crc = C.c_int(0)
        f = self.fn("decode", C.c_long, [C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t, i32p])
        r = f(symbols.ctypes.data, symbols.size, out.ctypes.data, cap, C.byref(crc))
        return int(r), out[:max(int(r), 0)].copy(), bool(crc.value)

    def window(self, sf, hann=1):
        assert not self.is_ref
        out = np.zeros(1 << sf, dtype=np.float32)
        self.fn("window", None, [C.c_uint, C.c_int, C.c_void_p])(sf, hann, out.ctypes.data)
        return out

    # ---- upstream coding chain (oracle restatement only) ------------------------------------
    def chain_cfg(self, sf, cr=1, explicit_header=1, crc=1, ppm=0, whitening=1, error_check=0, crc_check=1, out_header=0,
                  data_length=0):
        return ChainCfg(sf, ppm, cr, explicit_header, crc, whitening, error_check, crc_check, out_header, data_length)

    def encode_chain(self, cfg, payload):
        assert not self.is_ref
        payload = np.ascontiguousarray(payload, dtype=np.uint8)
        n = self.fn("chain_num_symbols", C.c_size_t, [C.c_void_p, C.c_size_t])(C.addressof(cfg), payload.size)
        out = np.zeros(n, dtype=np.uint16)
        r = self.fn("encode_chain", C.c_long, [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p])(
            C.addressof(cfg), payload.ctypes.data, payload.size, out.ctypes.data)
        assert r == n, (r, n)
        return out

    def decode_chain(self, cfg, symbols, cap=600):
        assert not self.is_ref
        symbols = np.ascontiguousarray(symbols, dtype=np.uint16)
        out = np.zeros(cap, dtype=np.uint8)
        fl = C.c_int(0)
        r = self.fn("decode_chain", C.c_long, [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t, i32p])(
            C.addressof(cfg), symbols.ctypes.data, symbols.size, out.ctypes.data, cap, C.byref(fl))
        return int(r), out[:max(int(r), 0)].copy(), fl.value

    # ---- AWGN Monte-Carlo spot check ------------------------------------------------------------
    def philox(self, ctr, key):
        c = (C.c_uint32 * 4)(*ctr); k = (C.c_uint32 * 2)(*key); o = (C.c_uint32 * 4)()
        self.fn("philox4x32", None, [C.c_void_p, C.c_void_p, C.c_void_p])(c, k, o)
        return list(o)

    def mc_payload(self, seed, cell, packet, n):
        out = np.zeros(n, np.uint8)
        self.fn("mc_payload", None, [C.c_uint64, C.c_uint32, C.c_uint64, C.c_size_t, C.c_void_p])(seed, cell, packet, n,
                                                                                                 out.ctypes.data)
        return out

    def mc_noise(self, seed, cell, packet, n, std):
        f = self.fn("mc_noise_sample", None, [C.c_uint64, C.c_uint32, C.c_uint64, C.c_uint32, C.c_float,
                                              C.POINTER(C.c_float), C.POINTER(C.c_float)])
        out = np.zeros(n, np.complex64)
        re, im = C.c_float(0), C.c_float(0)
        for i in range(n):
            f(seed, cell, packet, i, std, C.byref(re), C.byref(im))
            out[i] = re.value + 1j * im.value
        return out

    def awgn_roundtrip(self, cfg, noise_std, seed, cell, packet, payload_len):
        """-> dict(tx_payload, tx_sym, rx_sym, rx_payload, length, flags) for one packet of the sweep"""
        nsym = self.fn("chain_num_symbols", C.c_size_t, [C.c_void_p, C.c_size_t])(C.addressof(cfg), payload_len)
        txp, rxp = np.zeros(max(payload_len, 1), np.uint8), np.zeros(600, np.uint8)
        txs, rxs = np.zeros(nsym, np.uint16), np.zeros(nsym, np.uint16)
        fl = C.c_int(0)
        f = self.fn("awgn_roundtrip", C.c_long, [C.c_void_p, C.c_float, C.c_uint64, C.c_uint32, C.c_uint64, C.c_size_t,
                                                 C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, i32p])
        r = f(C.addressof(cfg), noise_std, seed, cell, packet, payload_len, txp.ctypes.data, txs.ctypes.data,
              rxs.ctypes.data, rxp.ctypes.data, rxp.size, C.byref(fl))
        return dict(tx_payload=txp[:payload_len], tx_sym=txs, rx_sym=rxs, rx_payload=rxp[:max(int(r), 0)].copy(),
                    length=int(r), flags=fl.value)

    # ---- upstream streaming receiver (oracle restatement only) ----------------------------------
    def stream_demod(self, iq, sf, sync=0x12, thresh=-30.0, mtu=256, max_frames=16, osr=1, bw_k=1):
        assert not self.is_ref
        iq = np.ascontiguousarray(iq, dtype=np.complex64)
        start = np.zeros(max_frames, np.uint32)
        err = np.zeros(max_frames, np.int32)
        nsym = np.zeros(max_frames, np.int32)
        sym = np.zeros((max_frames, mtu), np.int16)
        end = C.c_uint32(0)
        f = self.fn("stream_demod_ex", C.c_long, [C.c_uint, C.c_uint, C.c_uint, C.c_uint, C.c_float, C.c_uint, C.c_void_p, C.c_size_t,
                                                  C.c_size_t, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(C.c_uint32)])
        n = f(sf, osr, bw_k, sync, thresh, mtu, iq.ctypes.data, iq.size, max_frames, start.ctypes.data, err.ctypes.data,
              nsym.ctypes.data, sym.ctypes.data, C.byref(end))
        return int(n), start, err, nsym, sym, end.value

    def stream_tables(self, sf, bw_k=1):
        assert not self.is_ref
        N = 1 << sf
        up, down, fine = np.zeros(N, np.complex64), np.zeros(N, np.complex64), np.zeros(N * 128, np.complex64)
        self.fn("stream_tables_bw", None, [C.c_uint, C.c_uint, C.c_void_p, C.c_void_p, C.c_void_p])(sf, bw_k, up.ctypes.data,
                                                                                                  down.ctypes.data, fine.ctypes.data)
        return up, down, fine

    def frame_modulate(self, symbols, sf, ovs=1, sync=0x12, ampl=0.3, padding=1, bw_k=1):
        assert not self.is_ref
        symbols = np.ascontiguousarray(symbols, dtype=np.uint16)
        n = self.fn("frame_samples", C.c_size_t, [C.c_uint, C.c_uint, C.c_size_t, C.c_uint])(sf, ovs, symbols.size, padding)
        out = np.zeros(n, np.complex64)
        f = self.fn("frame_modulate_bw", C.c_size_t, [C.c_void_p, C.c_size_t, C.c_void_p, C.c_uint, C.c_uint, C.c_uint,
                                                      C.c_float, C.c_uint, C.c_uint])
        got = f(symbols.ctypes.data, symbols.size, out.ctypes.data, sf, ovs, sync, ampl, padding, bw_k)
        assert got == n
        return out

    def stream_frame(self, symbols, sf, sync=0x12, ampl=1.0, pad_symbols=2, cfo_bins=0.0):
        """Upstream frame format (external/LoRa-SDR/LoRaMod.cpp:141-215): 10 upchirps, sync word0 / word1 chirps at
        (nibble * 8), 2 1/4 downchirps, data chirps, zero padding -- synthesised in float64 on the chirp the
        upstream receiver's own tables describe (LoRaDemod.cpp:96-106), so a clean frame lands on bin 0.
        TEST INPUT ONLY: parity is CUDA == oracle on whatever IQ this produces."""
        N = 1 << sf
        i = np.arange(N)
        acc = np.cumsum(-np.pi + 2 * np.pi * i / N)
        upc = np.exp(1j * acc)
        parts = [upc] * 10
        parts.append(upc * np.exp(2j * np.pi * ((sync >> 4) * 8) * i / N))
        parts.append(upc * np.exp(2j * np.pi * ((sync & 0xf) * 8) * i / N))
        parts += [np.conj(upc), np.conj(upc), np.conj(upc)[: N // 4]]
        for sym in symbols:
            parts.append(upc * np.exp(2j * np.pi * int(sym) * i / N))
        x = ampl * np.concatenate(parts)
        if cfo_bins:
            x = x * np.exp(2j * np.pi * cfo_bins * np.arange(x.size) / N)
        return np.concatenate([x, np.zeros(pad_symbols * N)]).astype(np.complex64)

    # ---- timed batch ---------------------------------------------------------------
    def demod_batch_mt(self, iq, n_packets, spp, sf, osr=1, window=0, threads=1):
        iq = np.ascontiguousarray(iq, dtype=np.complex64)
        nsym = spp // ((1 << sf) * osr)
        stride = max(nsym - 2, 1)
        out = np.zeros(n_packets * stride, dtype=np.uint16)
        sync = np.zeros(n_packets, dtype=np.uint8)
        f = self.fn("demod_batch_mt", C.c_double,
                    [C.c_uint, C.c_int, C.c_uint, C.c_void_p, C.c_size_t, C.c_size_t, C.c_void_p,
                     C.c_size_t, C.c_void_p, C.c_int])
        sec = f(sf, window, osr, iq.ctypes.data, n_packets, spp, out.ctypes.data, stride,
                sync.ctypes.data, threads)
        return sec, out.reshape(n_packets, stride), sync


class RefWorkspace:
    """High-level workspace API of the reference (src/phy/phy.cpp) through ref_shim."""

    def __init__(self, ref, sf, bw_hz=125000, cr=1, osr=1, window=0, sync=0x12):
        self.l, self.sf, self.osr = ref.lib, sf, osr
        rc = C.c_int(0)
        self.l.ref_ws_create.restype = C.c_void_p
        self.l.ref_ws_create.argtypes = [C.c_uint, C.c_uint, C.c_uint, C.c_uint, C.c_int, C.c_uint8, i32p]
        self.h = C.c_void_p(self.l.ref_ws_create(sf, bw_hz, cr, osr, window, sync, C.byref(rc)))
        self.rc = rc.value

    def close(self):
        if self.h:
            self.l.ref_ws_destroy.argtypes = [C.c_void_p]
            self.l.ref_ws_destroy(self.h)
            self.h = None

    def demodulate(self, iq, cap=None):
        iq = np.ascontiguousarray(iq, dtype=np.complex64)
        cap = iq.size if cap is None else cap
        out = np.zeros(max(cap, 1), dtype=np.uint16)
        f = self.l.ref_ws_demodulate
        f.restype, f.argtypes = C.c_long, [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t]
        r = f(self.h, iq.ctypes.data, iq.size, out.ctypes.data, cap)
        return int(r), out[:max(int(r), 0)].copy()

    def estimate_offsets(self, iq):
        iq = np.ascontiguousarray(iq, dtype=np.complex64)
        f = self.l.ref_ws_estimate_offsets
        f.restype, f.argtypes = None, [C.c_void_p, C.c_void_p, C.c_size_t]
        f(self.h, iq.ctypes.data, iq.size)

    def compensate_offsets(self, iq, cfo, to):
        iq = np.array(iq, dtype=np.complex64, copy=True)
        g = self.l.ref_ws_set_metrics
        g.restype, g.argtypes = None, [C.c_void_p, C.c_float, C.c_float]
        g(self.h, cfo, to)
        f = self.l.ref_ws_compensate_offsets
        f.restype, f.argtypes = None, [C.c_void_p, C.c_void_p, C.c_size_t]
        f(self.h, iq.ctypes.data, iq.size)
        return iq

    def decode(self, symbols, cap=None):
        symbols = np.ascontiguousarray(symbols, dtype=np.uint16)
        cap = symbols.size if cap is None else cap
        out = np.zeros(symbols.size // 2 + 8, dtype=np.uint8)
        f = self.l.ref_ws_decode
        f.restype, f.argtypes = C.c_long, [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t]
        r = f(self.h, symbols.ctypes.data, symbols.size, out.ctypes.data, cap)
        return int(r), out[:max(int(r), 0)].copy()

    def metrics(self):
        crc, cfo, to, sync = C.c_int(), C.c_float(), C.c_float(), C.c_uint8()
        f = self.l.ref_ws_get_metrics
        f.restype, f.argtypes = None, [C.c_void_p, i32p, f32p, f32p, C.POINTER(C.c_uint8)]
        f(self.h, C.byref(crc), C.byref(cfo), C.byref(to), C.byref(sync))
        return bool(crc.value), np.float32(cfo.value), np.float32(to.value), sync.value


def build_checkers():
    subprocess.run(["make", "-s", "-C", os.path.join(ROOT, "oracle"), "oracle", "CC=gcc"], check=True)
    if os.path.isdir("/root/reference/src/phy"):
        subprocess.run(["make", "-s", "-C", os.path.join(ROOT, "oracle"), "ref", "CXX=g++"], check=True)


def load_oracle():
    if not os.path.exists(ORACLE_SO):
        build_checkers()
    return CpuPhy(C.CDLL(ORACLE_SO), "oracle_")


def load_ref():
    if not os.path.exists(REF_SO):
        if os.path.isdir("/root/reference/src/phy"):
            build_checkers()
        else:
            return None
    return CpuPhy(C.CDLL(REF_SO), "ref_")
