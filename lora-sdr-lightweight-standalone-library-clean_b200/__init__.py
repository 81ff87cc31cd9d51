"""B200-native batched LoRa PHY -- Python harness over the C ABI (include/lora_phy_cuda.h).

The product is ``liblora_phy_cuda.so`` (hand-written sm_100a kernels behind a plain C ABI).
This module is harness plumbing only: it loads the library with ctypes and passes raw device
pointers of torch tensors.  There is no CPU or PyTorch fallback: if the library is missing or
no B200 is present, construction raises.
"""
import ctypes as C
import errno
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(_HERE)
LIB_PATH = os.environ.get("LORA_PHY_LIB") or os.path.join(_HERE, "liblora_phy_cuda.so")   # override: A/B of build variants
HEADER = os.path.join(ROOT, "include", "lora_phy_cuda.h")

F_DECHIRP = 0x1
F_NO_SCRATCH = 0x2
F_ONLY_PREPASS = 0x10
F_ONLY_SYMBOLS = 0x20
F_TWO_KERNEL = 0x80
F_EXACT = 0x100
STATIC_MAX_CHIRPS = 1 << 20        # LORA_PHY_CUDA_STATIC_MAX_CHIRPS


class Cfg(C.Structure):
    _fields_ = [("sf", C.c_uint32), ("bw_hz", C.c_uint32), ("osr", C.c_uint32), ("window", C.c_uint32),
                ("flags", C.c_uint32), ("sync_word", C.c_uint32), ("amplitude", C.c_float),
                ("reserved", C.c_uint32)]


def make_cfg(sf, bw_hz=125000, osr=1, window=0, flags=0, sync_word=0x12, amplitude=1.0):
    return Cfg(sf, bw_hz, osr, window, flags, sync_word, amplitude, 0)


class ChainCfg(C.Structure):
    _fields_ = [(n, C.c_uint32) for n in ("sf", "ppm", "cr", "explicit_header", "crc", "whitening", "error_check",
                                          "crc_check", "out_header", "data_length")]


def make_chain_cfg(sf, cr=1, explicit_header=1, crc=1, ppm=0, whitening=1, error_check=0, crc_check=1, out_header=0,
                   data_length=0):
    return ChainCfg(sf, ppm, cr, explicit_header, crc, whitening, error_check, crc_check, out_header, data_length)


def build(verbose=False):
    """Compile csrc/*.cu for sm_100a into liblora_phy_cuda.so (nvcc cross-compiles without a GPU)."""
    cmd = ["make", "-j4", "-C", os.path.join(_HERE, "csrc")]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if verbose or r.returncode:
        print(r.stdout[-4000:], r.stderr[-4000:])
    if r.returncode:
        raise RuntimeError("nvcc build of liblora_phy_cuda.so failed")
    return LIB_PATH


_lib = None

_vp, _sz, _u32, _i32, _f32 = C.c_void_p, C.c_size_t, C.c_uint32, C.c_int, C.c_float
_cfgp = C.POINTER(Cfg)

# name -> (restype, argtypes); mirrors include/lora_phy_cuda.h one to one
SIGNATURES = {
    "lora_phy_cuda_abi_version": (C.c_int, []),
    "lora_phy_cuda_create": (C.c_int, [C.POINTER(_vp), C.c_int]),
    "lora_phy_cuda_destroy": (None, [_vp]),
    "lora_phy_cuda_last_error": (C.c_char_p, [_vp]),
    "lora_phy_cuda_device": (C.c_int, [_vp]),
    "lora_phy_cuda_launch_count": (C.c_uint64, [_vp]),
    "lora_phy_cuda_set_stream": (C.c_int, [_vp, _vp]),
    "lora_phy_cuda_synchronize": (C.c_int, [_vp]),
    "lora_phy_cuda_cert_stats": (C.c_int, [_vp, C.POINTER(C.c_uint64)]),
    "lora_phy_cuda_cert_stats_ex": (C.c_int, [_vp, C.POINTER(C.c_uint64), _sz]),
    "lora_phy_cuda_demodulate_batch": (C.c_int, [_vp, _cfgp, _vp, _sz, _vp, _sz, _sz, _vp, _sz, _vp, _vp, _vp, _vp]),
    "lora_phy_cuda_demodulate_batch_host": (C.c_int, [_vp, _cfgp, _vp, _sz, _vp, _sz, _sz, _vp, _sz, _vp, _vp, _vp, _vp]),
    "lora_phy_cuda_estimate_offsets_batch": (C.c_int, [_vp, _cfgp, _vp, _sz, _vp, _sz, _sz, _vp, _vp]),
    "lora_phy_cuda_compensate_offsets_batch": (C.c_int, [_vp, _cfgp, _vp, _sz, _sz, _sz, _vp, _vp]),
    "lora_phy_cuda_estimate_offsets_batch_host": (C.c_int, [_vp, _cfgp, _vp, _sz, _sz, _sz, _vp, _vp]),
    "lora_phy_cuda_compensate_offsets_batch_host": (C.c_int, [_vp, _cfgp, _vp, _sz, _sz, _sz, _vp, _vp]),
    "lora_phy_cuda_detect_batch": (C.c_int, [_vp, _u32, _vp, _sz, _vp, _vp, _vp, _vp, _vp]),
    "lora_phy_cuda_modulate_batch": (C.c_long, [_vp, _cfgp, _vp, _sz, _sz, _sz, _vp, _sz]),
    "lora_phy_cuda_modulate_batch_host": (C.c_long, [_vp, _cfgp, _vp, _sz, _sz, _sz, _vp, _sz]),
    "lora_phy_cuda_modulate_batch_static": (C.c_long, [_vp, _cfgp, _vp, _sz, _sz, _sz, _vp, _sz]),
    "lora_phy_cuda_tx_batch": (C.c_long, [_vp, _cfgp, C.POINTER(ChainCfg), _vp, _sz, _sz, _sz, _vp, _sz]),
    "lora_phy_cuda_tx_batch_host": (C.c_long, [_vp, _cfgp, C.POINTER(ChainCfg), _vp, _sz, _sz, _sz, _vp, _sz]),
    "lora_phy_cuda_rx_batch": (C.c_int, [_vp, _cfgp, C.POINTER(ChainCfg), _vp, _sz, _sz, _sz, _vp, _sz, _vp, _vp, _vp, _vp, _vp]),
    "lora_phy_cuda_rx_batch_host": (C.c_int, [_vp, _cfgp, C.POINTER(ChainCfg), _vp, _sz, _sz, _sz, _vp, _sz, _vp, _vp, _vp, _vp,
                                              _vp]),
    "lora_phy_cuda_awgn_roundtrip": (C.c_int, [_vp, C.POINTER(ChainCfg), C.c_float, C.c_uint64, C.c_uint32, C.c_uint64, _sz,
                                               _sz, _vp, _vp, _vp, _sz, _vp, _sz, _vp]),
    "lora_phy_cuda_frame_samples": (C.c_long, [C.c_uint32, C.c_uint32, _sz, C.c_uint32]),
    "lora_phy_cuda_modulate_frames_batch": (C.c_long, [_vp, _cfgp, C.c_uint32, _vp, _sz, _sz, _sz, _vp, _sz]),
    "lora_phy_cuda_dechirp_batch": (C.c_int, [_vp, _cfgp, _vp, _sz, _sz, _sz]),
    "lora_phy_cuda_encode_batch": (C.c_long, [_vp, _vp, _sz, _sz, _sz, _vp, _sz]),
    "lora_phy_cuda_decode_batch": (C.c_long, [_vp, _vp, _sz, _vp, _sz, _sz, _vp, _sz, _vp]),
    "lora_phy_cuda_encode_batch_host": (C.c_long, [_vp, _vp, _sz, _sz, _sz, _vp, _sz]),
    "lora_phy_cuda_decode_batch_host": (C.c_long, [_vp, _vp, _sz, _vp, _sz, _sz, _vp, _sz, _vp]),
    "lora_phy_cuda_detect_stream": (C.c_int, [_vp, _u32, _u32, _f32, _u32, _vp, _sz, _vp, _sz, _sz, _sz, _vp, _vp, _vp, _vp, _vp, _vp]),
    "lora_phy_cuda_detect_streams_mixed": (C.c_int, [_vp, _vp, _u32, _f32, _u32, _vp, _sz, _vp, _sz, _sz, _sz, _vp, _vp, _vp, _vp,
                                                     _vp, _vp]),
    "lora_phy_cuda_detect_streams_ex": (C.c_int, [_vp, _vp, _vp, _vp, _u32, _f32, _u32, _vp, _sz, _vp, _sz, _sz, _sz, _vp, _vp, _vp, _vp,
                                                  _vp, _vp, _vp, _vp]),
    "lora_phy_cuda_chain_symbols": (C.c_long, [_vp, _sz]),
    "lora_phy_cuda_encode_chain_batch": (C.c_long, [_vp, _vp, _vp, _sz, _sz, _sz, _vp, _sz]),
    "lora_phy_cuda_decode_chain_batch": (C.c_int, [_vp, _vp, _vp, _sz, _vp, _sz, _sz, _vp, _sz, _vp, _vp]),
    "lora_phy_cuda_copy_table": (C.c_int, [_vp, _u32, _u32, C.c_int, _vp]),
}


def load_library():
    """dlopen the C-ABI library and bind every symbol the header declares.  No compute happens."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(f"{LIB_PATH} is missing: run __graft_entry__.build() (there is no fallback path)")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)          # AttributeError if the library does not export it
        fn.restype, fn.argtypes = res, args
    _lib = lib
    return lib


class LoraPhyError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"{errno.errorcode.get(-code, code)}: {msg}")
        self.code = code


def _ptr(t):
    return None if t is None else C.c_void_p(t.data_ptr())


class LoraPhyCuda:
    """One context per GPU.  All tensor arguments are torch CUDA tensors on that GPU."""

    def __init__(self, device=0):
        import torch
        self.torch = torch
        self.lib = load_library()
        h = _vp()
        rc = self.lib.lora_phy_cuda_create(C.byref(h), int(device))
        if rc:
            raise LoraPhyError(rc, "lora_phy_cuda_create failed (a B200 / sm_100 device is required; no fallback)")
        self.h = h
        self.device = torch.device("cuda", int(device))
        # enqueue on torch's current stream so tensor allocation / fill / copies order with our kernels
        with torch.cuda.device(self.device):
            self.set_stream(torch.cuda.current_stream().cuda_stream)

    def close(self):
        if getattr(self, "h", None):
            self.lib.lora_phy_cuda_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc):
        if rc < 0:
            raise LoraPhyError(rc, self.lib.lora_phy_cuda_last_error(self.h).decode())
        return rc

    def launch_count(self):
        return int(self.lib.lora_phy_cuda_launch_count(self.h))

    def synchronize(self):
        self._check(self.lib.lora_phy_cuda_synchronize(self.h))

    def cert_stats(self):
        """(windows, fallback_windows, second_estimates, rewalked_packets, pruned_windows, lean_flagged_packets) of the
        certified demodulator so far."""
        out = (C.c_uint64 * 6)()
        self._check(self.lib.lora_phy_cuda_cert_stats_ex(self.h, out, 6))
        return tuple(int(x) for x in out)

    def set_stream(self, stream_handle):
        self._check(self.lib.lora_phy_cuda_set_stream(self.h, _vp(stream_handle)))

    # ---- demodulator ---------------------------------------------------------------------
    def alloc_demod_outputs(self, n_packets, symbol_stride):
        t = self.torch
        return dict(symbols=t.zeros((n_packets, max(symbol_stride, 1)), dtype=t.uint16, device=self.device),
                    sync=t.zeros(n_packets, dtype=t.uint8, device=self.device),
                    cfo=t.zeros(n_packets, dtype=t.float32, device=self.device),
                    time_offset=t.zeros(n_packets, dtype=t.float32, device=self.device),
                    status=t.zeros(n_packets, dtype=t.int32, device=self.device))

    def demodulate_batch(self, cfg, iq, samples_per_packet=None, sample_counts=None, out=None, raw_rc=False):
        """iq: complex64 [P, stride].  Returns dict of output tensors (asynchronous)."""
        P, stride = iq.shape
        spp = stride if samples_per_packet is None else samples_per_packet
        step = (1 << cfg.sf) * cfg.osr
        total = spp // step
        sym_stride = max(total - 2, 1) if total >= 2 else max(total, 1)
        if out is None:
            out = self.alloc_demod_outputs(P, sym_stride)
        rc = self.lib.lora_phy_cuda_demodulate_batch(
            self.h, C.byref(cfg), _ptr(iq), stride, _ptr(sample_counts), spp, P, _ptr(out["symbols"]),
            out["symbols"].shape[1], _ptr(out["sync"]), _ptr(out["cfo"]), _ptr(out["time_offset"]), _ptr(out["status"]))
        if raw_rc:
            return rc, out
        self._check(rc)
        return out

    def demodulate_batch_host(self, cfg, iq_np, samples_per_packet=None, sample_counts=None, symbol_stride=None):
        """numpy in / numpy out through the host-buffer entry point (copies inside the call)."""
        import numpy as np
        iq_np = np.ascontiguousarray(iq_np, dtype=np.complex64)
        P, stride = iq_np.shape
        spp = stride if samples_per_packet is None else samples_per_packet
        step = (1 << cfg.sf) * cfg.osr
        total = spp // step
        if symbol_stride is None:
            symbol_stride = max(total - 2, 1) if total >= 2 else max(total, 1)
        sym = np.zeros((P, symbol_stride), dtype=np.uint16)
        sync = np.zeros(P, dtype=np.uint8)
        cfo = np.zeros(P, dtype=np.float32)
        to = np.zeros(P, dtype=np.float32)
        st = np.zeros(P, dtype=np.int32)
        cnt = None if sample_counts is None else np.ascontiguousarray(sample_counts, dtype=np.uint32)
        rc = self.lib.lora_phy_cuda_demodulate_batch_host(
            self.h, C.byref(cfg), iq_np.ctypes.data, stride, None if cnt is None else cnt.ctypes.data, spp, P,
            sym.ctypes.data, symbol_stride, sync.ctypes.data, cfo.ctypes.data, to.ctypes.data, st.ctypes.data)
        return rc, dict(symbols=sym, sync=sync, cfo=cfo, time_offset=to, status=st)

    def estimate_offsets_batch(self, cfg, iq, samples_per_packet=None):
        t = self.torch
        P, stride = iq.shape
        spp = stride if samples_per_packet is None else samples_per_packet
        cfo = t.zeros(P, dtype=t.float32, device=self.device)
        to = t.zeros(P, dtype=t.float32, device=self.device)
        self._check(self.lib.lora_phy_cuda_estimate_offsets_batch(self.h, C.byref(cfg), _ptr(iq), stride, None, spp, P,
                                                                  _ptr(cfo), _ptr(to)))
        return cfo, to

    def compensate_offsets_batch(self, cfg, iq, cfo, time_offset):
        P, stride = iq.shape
        self._check(self.lib.lora_phy_cuda_compensate_offsets_batch(self.h, C.byref(cfg), _ptr(iq), stride, stride, P,
                                                                    _ptr(cfo), _ptr(time_offset)))
        return iq

    def detect_batch(self, sf, windows, want_spectrum=True):
        t = self.torch
        W = windows.shape[0]
        spec = t.zeros((W, 1 << sf), dtype=t.complex64, device=self.device) if want_spectrum else None
        idx = t.zeros(W, dtype=t.int32, device=self.device)
        p = t.zeros(W, dtype=t.float32, device=self.device)
        pav = t.zeros(W, dtype=t.float32, device=self.device)
        fi = t.zeros(W, dtype=t.float32, device=self.device)
        self._check(self.lib.lora_phy_cuda_detect_batch(self.h, sf, _ptr(windows), W, _ptr(spec), _ptr(idx), _ptr(p),
                                                        _ptr(pav), _ptr(fi)))
        return dict(spectrum=spec, index=idx, power=p, power_avg=pav, findex=fi)

    # ---- modulator / dechirp ---------------------------------------------------------------
    def modulate_batch(self, cfg, symbols, iq=None):
        t = self.torch
        P, S = symbols.shape
        n = (S + 2) * (1 << cfg.sf) * cfg.osr
        if iq is None:
            iq = t.empty((P, n), dtype=t.complex64, device=self.device)
        self._check(self.lib.lora_phy_cuda_modulate_batch(self.h, C.byref(cfg), _ptr(symbols), symbols.stride(0), S, P,
                                                          _ptr(iq), iq.stride(0)))
        return iq

    def modulate_batch_static(self, cfg, symbols, iq):
        """lora_phy_cuda_modulate_batch_static: allocation-free (static device scratch, at most STATIC_MAX_CHIRPS chirps
        per call); the caller passes the output.  No counterpart in the reference."""
        P, S = symbols.shape
        self._check(self.lib.lora_phy_cuda_modulate_batch_static(self.h, C.byref(cfg), _ptr(symbols), symbols.stride(0), S, P,
                                                                 _ptr(iq), iq.stride(0)))
        return iq

    def modulate_frames_batch(self, cfg, symbols, padding=1, iq=None):
        """Upstream frame format (preamble, sync word, 2 1/4 down-chirps, data, padding), one frame per row."""
        t = self.torch
        P, S = symbols.shape
        n = int(self.lib.lora_phy_cuda_frame_samples(cfg.sf, cfg.osr, S, padding))
        if n < 0:
            raise LoraPhyError(n, "bad frame geometry")
        if iq is None:
            iq = t.empty((P, n), dtype=t.complex64, device=self.device)
        self._check(self.lib.lora_phy_cuda_modulate_frames_batch(self.h, C.byref(cfg), padding, _ptr(symbols),
                                                                 symbols.stride(0), S, P, _ptr(iq), iq.stride(0)))
        return iq

    def dechirp_batch(self, cfg, iq):
        P, stride = iq.shape
        self._check(self.lib.lora_phy_cuda_dechirp_batch(self.h, C.byref(cfg), _ptr(iq), iq.stride(0), stride, P))
        return iq

    # ---- shipped coder ----------------------------------------------------------------------
    def encode_batch(self, payload):
        t = self.torch
        P, L = payload.shape
        sym = t.zeros((P, 2 * L), dtype=t.uint16, device=self.device)
        self._check(self.lib.lora_phy_cuda_encode_batch(self.h, _ptr(payload), payload.stride(0), L, P, _ptr(sym), 2 * L))
        return sym

    def decode_batch(self, symbols, counts=None, n_symbols=None):
        t = self.torch
        P, S = symbols.shape
        S = S if n_symbols is None else n_symbols
        pay = t.zeros((P, max(S // 2, 1)), dtype=t.uint8, device=self.device)
        crc = t.zeros(P, dtype=t.uint8, device=self.device)
        rc = self._check(self.lib.lora_phy_cuda_decode_batch(self.h, _ptr(symbols), symbols.stride(0), _ptr(counts), S, P,
                                                             _ptr(pay), pay.stride(0), _ptr(crc)))
        return rc, pay, crc

    # ---- streaming front end ---------------------------------------------------------------------
    def detect_stream(self, sf, iq, sync_word=0x12, thresh_db=-30.0, mtu=256, max_frames=16, sample_counts=None, bw_scale=None,
                      osr=None, cfo=None, time_offset=None):
        """iq: complex64 [n_streams, n_samples] continuous IQ.  Returns dict of per-stream frame reports.
        sf / bw_scale / osr: one value, or a sequence with one per stream (mixed batch, cfg 4); cfo / time_offset:
        optional per-stream float tensors applied as compensate_offsets inside the window load."""
        import numpy as np
        t = self.torch
        S, n = iq.shape
        out = dict(frame_start=t.zeros((S, max_frames), dtype=t.int32, device=self.device),
                   frame_err=t.zeros((S, max_frames), dtype=t.int32, device=self.device),
                   frame_nsym=t.zeros((S, max_frames), dtype=t.int32, device=self.device),
                   symbols=t.zeros((S, max_frames, mtu), dtype=t.int16, device=self.device),
                   n_frames=t.zeros(S, dtype=t.int32, device=self.device),
                   end_pos=t.zeros(S, dtype=t.int32, device=self.device))
        tail = (_ptr(out["frame_start"]), _ptr(out["frame_err"]), _ptr(out["frame_nsym"]), _ptr(out["symbols"]),
                _ptr(out["n_frames"]), _ptr(out["end_pos"]))
        plain = bw_scale is None and osr is None and cfo is None and time_offset is None
        if isinstance(sf, int) and plain:
            self._check(self.lib.lora_phy_cuda_detect_stream(self.h, sf, sync_word, thresh_db, mtu, _ptr(iq), iq.stride(0),
                                                             _ptr(sample_counts), n, S, max_frames, *tail))
            return out

        def per_stream(v, default):
            if v is None:
                v = default
            a = np.ascontiguousarray(np.broadcast_to(np.asarray(v, dtype=np.uint8), (S,)))
            return a
        sfs = per_stream(sf, 7)
        if plain:
            self._check(self.lib.lora_phy_cuda_detect_streams_mixed(self.h, sfs.ctypes.data, sync_word, thresh_db, mtu, _ptr(iq),
                                                                    iq.stride(0), _ptr(sample_counts), n, S, max_frames, *tail))
            return out
        bws, osrs = per_stream(bw_scale, 1), per_stream(osr, 1)
        self._check(self.lib.lora_phy_cuda_detect_streams_ex(self.h, sfs.ctypes.data, bws.ctypes.data, osrs.ctypes.data, sync_word,
                                                             thresh_db, mtu, _ptr(iq), iq.stride(0), _ptr(sample_counts), n, S,
                                                             max_frames, _ptr(cfo), _ptr(time_offset), *tail))
        return out

    # ---- full coding chain --------------------------------------------------------------------
    def chain_symbols(self, cc, payload_len):
        return int(self.lib.lora_phy_cuda_chain_symbols(C.byref(cc), payload_len))

    def encode_chain_batch(self, cc, payload):
        t = self.torch
        P, L = payload.shape
        nsym = self.chain_symbols(cc, L)
        if nsym < 0:
            raise LoraPhyError(nsym, "bad chain configuration")
        sym = t.zeros((P, nsym), dtype=t.uint16, device=self.device)
        self._check(self.lib.lora_phy_cuda_encode_chain_batch(self.h, C.byref(cc), _ptr(payload), payload.stride(0), L, P,
                                                              _ptr(sym), nsym))
        return sym

    def decode_chain_batch(self, cc, symbols, counts=None, payload_stride=260):
        t = self.torch
        P, S = symbols.shape
        pay = t.zeros((P, payload_stride), dtype=t.uint8, device=self.device)
        ln = t.zeros(P, dtype=t.int32, device=self.device)
        fl = t.zeros(P, dtype=t.uint8, device=self.device)
        self._check(self.lib.lora_phy_cuda_decode_chain_batch(self.h, C.byref(cc), _ptr(symbols), symbols.stride(0),
                                                              _ptr(counts), S, P, _ptr(pay), payload_stride, _ptr(ln), _ptr(fl)))
        return pay, ln, fl

    # ---- whole-modem pipelines ---------------------------------------------------------------
    def tx_batch(self, cfg, cc, payload):
        t = self.torch
        P, L = payload.shape
        n = (self.chain_symbols(cc, L) + 2) * (1 << cfg.sf) * cfg.osr
        iq = t.empty((P, n), dtype=t.complex64, device=self.device)
        self._check(self.lib.lora_phy_cuda_tx_batch(self.h, C.byref(cfg), C.byref(cc), _ptr(payload), payload.stride(0), L, P,
                                                    _ptr(iq), iq.stride(0)))
        return iq

    def rx_batch(self, cfg, cc, iq, payload_stride=260):
        """iq is dechirped in place.  -> dict(payload, len, flags, sync, cfo, time_offset)"""
        t = self.torch
        P, n = iq.shape
        out = dict(payload=t.zeros((P, payload_stride), dtype=t.uint8, device=self.device),
                   len=t.zeros(P, dtype=t.int32, device=self.device), flags=t.zeros(P, dtype=t.uint8, device=self.device),
                   sync=t.zeros(P, dtype=t.uint8, device=self.device), cfo=t.zeros(P, dtype=t.float32, device=self.device),
                   time_offset=t.zeros(P, dtype=t.float32, device=self.device))
        self._check(self.lib.lora_phy_cuda_rx_batch(self.h, C.byref(cfg), C.byref(cc), _ptr(iq), iq.stride(0), n, P,
                                                    _ptr(out["payload"]), payload_stride, _ptr(out["len"]), _ptr(out["flags"]),
                                                    _ptr(out["sync"]), _ptr(out["cfo"]), _ptr(out["time_offset"])))
        return out

    def awgn_roundtrip(self, cc, snr_db, n_packets, payload_len, seed=0xA76E, cell_id=0, first_packet=0, counts=None,
                       details=False, noise_std=None):
        """One cell of the Monte-Carlo sweep.  Returns (counts tensor[8] uint64-as-int64, details dict or None)."""
        import math
        t = self.torch
        if noise_std is None:
            noise_std = math.sqrt(10.0 ** (-snr_db / 10.0) / 2.0)
        if counts is None:
            counts = t.zeros(8, dtype=t.int64, device=self.device)
        det = None
        args = [None, 0, None, 0, None]
        ok = None
        if details:
            nsym = self.chain_symbols(cc, payload_len)
            det = dict(ok=t.zeros(n_packets, dtype=t.uint8, device=self.device),
                       rx_symbols=t.zeros((n_packets, nsym), dtype=t.uint16, device=self.device),
                       rx_payload=t.zeros((n_packets, max(payload_len, 1)), dtype=t.uint8, device=self.device),
                       rx_len=t.zeros(n_packets, dtype=t.int32, device=self.device))
            ok = det["ok"]
            args = [_ptr(det["rx_symbols"]), nsym, _ptr(det["rx_payload"]), max(payload_len, 1), _ptr(det["rx_len"])]
        self._check(self.lib.lora_phy_cuda_awgn_roundtrip(self.h, C.byref(cc), noise_std, seed, cell_id, first_packet,
                                                          n_packets, payload_len, _ptr(counts), _ptr(ok), *args))
        return counts, det

    def copy_table(self, sf, which, bw_hz=125000):
        import numpy as np
        N = 1 << sf
        out = np.zeros(N, dtype=np.float32 if which == 1 else np.complex64)
        self._check(self.lib.lora_phy_cuda_copy_table(self.h, sf, bw_hz, which, out.ctypes.data))
        return out


from .shard import packet_range, all_ranges, gather_results  # noqa: E402  (multi-GPU packet-range sharding)
