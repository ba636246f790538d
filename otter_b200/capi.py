"""ctypes binding of include/otter_b200.h (one function per C entry point, nothing more)."""
import ctypes as C
import os

import numpy as np

LIB_PATH = os.environ.get("OTTER_B200_LIB") or os.path.join(
    os.path.dirname(os.path.abspath(__file__)), "lib", "libotter_b200.so")  # env: kernel experiments

# Every symbol include/otter_b200.h declares; tests assert the library exports all of them.
SYMBOLS = [
    "otgpu_abi_version", "otgpu_device_count", "otgpu_create", "otgpu_destroy", "otgpu_strerror",
    "otgpu_last_cuda_error", "otgpu_stream", "otgpu_get_stats", "otgpu_host_alloc",
    "otgpu_host_free", "otgpu_edit_scores", "otgpu_region_edit_scores",
    "otgpu_batch_upload_regions", "otgpu_batch_upload_pairs", "otgpu_batch_align",
    "otgpu_batch_download", "otgpu_synchronize", "otgpu_last_align_ms", "otgpu_affine_cigars",
    "otgpu_edit_cigars", "otgpu_last_chunk_count", "otgpu_first_window", "otgpu_affine_cigars_packed",
    "otgpu_edit_cigars_packed", "otgpu_extract_reads", "otgpu_kde_densities",
    "otgpu_region_edit_scores_hooked", "otgpu_region_edit_scores_extracted",
    "otgpu_edit_scores_static", "otgpu_ppoa_consensus",
]


class OtgpuError(RuntimeError):
    pass


class Pair(C.Structure):
    _fields_ = [("pattern", C.c_uint32), ("text", C.c_uint32), ("pattern_begin_free", C.c_int32),
                ("pattern_end_free", C.c_int32), ("text_begin_free", C.c_int32),
                ("text_end_free", C.c_int32)]


PAIR_DTYPE = np.dtype([("pattern", "<u4"), ("text", "<u4"), ("pbf", "<i4"), ("pef", "<i4"),
                       ("tbf", "<i4"), ("tef", "<i4")])


# otgpu_bam_read (include/otter_b200.h)
BAM_READ_DTYPE = np.dtype([("pos", "<i4"), ("l_qseq", "<i4"), ("n_cigar", "<u4"), ("reserved", "<u4"),
                           ("cigar_off", "<u8"), ("seq_off", "<u8"), ("start", "<i4"), ("end", "<i4")])
NT16 = "=ACMGRSVTWYHKDBN"


def encode_seq4(seq):
    """BAM 4-bit encoding of a base string (two bases per byte, high nibble first)."""
    table = np.full(256, 15, dtype=np.uint8)
    for i, ch in enumerate(NT16):
        table[ord(ch)] = i
        table[ord(ch.lower())] = i
    codes = table[np.frombuffer(seq.encode() if isinstance(seq, str) else bytes(seq), dtype=np.uint8)]
    if len(codes) % 2:
        codes = np.concatenate([codes, np.zeros(1, np.uint8)])
    return (codes[0::2] << 4 | codes[1::2]).astype(np.uint8)


class Stats(C.Structure):
    _fields_ = [("n_pairs", C.c_uint64), ("n_pairs_8bit", C.c_uint64), ("n_pairs_wide", C.c_uint64),
                ("kernel_launches", C.c_uint64)]


_lib = None


def lib():
    """Load the CUDA library; fail loudly if it has not been built (no fallback exists)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise OtgpuError(f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; "
                             "g.build()'` (nvcc, sm_100a). otter_b200 has no CPU fallback.")
        l = C.CDLL(LIB_PATH)
        vp, u64, u32, i32 = C.c_void_p, C.c_uint64, C.c_uint32, C.c_int
        l.otgpu_abi_version.restype = i32
        l.otgpu_device_count.restype = i32
        l.otgpu_create.argtypes = [i32, C.POINTER(vp)]
        l.otgpu_destroy.argtypes = [vp]
        l.otgpu_destroy.restype = None
        l.otgpu_strerror.argtypes = [i32]
        l.otgpu_strerror.restype = C.c_char_p
        l.otgpu_last_cuda_error.argtypes = [vp]
        l.otgpu_last_cuda_error.restype = C.c_char_p
        l.otgpu_stream.argtypes = [vp]
        l.otgpu_stream.restype = vp
        l.otgpu_get_stats.argtypes = [vp, C.POINTER(Stats)]
        l.otgpu_host_alloc.argtypes = [C.c_size_t, C.POINTER(vp)]
        l.otgpu_host_free.argtypes = [vp]
        l.otgpu_host_free.restype = None
        l.otgpu_edit_scores.argtypes = [vp, vp, u64, vp, vp, u32, vp, u64, vp]
        l.otgpu_region_edit_scores.argtypes = [vp, vp, u64, vp, vp, u32, vp, u32, vp, u64, vp]
        l.otgpu_region_edit_scores_hooked.argtypes = [vp, vp, u64, vp, vp, u32, vp, u32, vp, u64, vp, vp]
        l.otgpu_region_edit_scores_extracted.argtypes = [vp, vp, u32, vp, u32, vp, u64, vp]
        l.otgpu_edit_scores_static.argtypes = [i32, vp, vp, u64, vp, vp, u32, vp, u64, vp]
        l.otgpu_ppoa_consensus.argtypes = [vp, vp, u64, vp, vp, u32, vp, u32, vp, u32, vp, u64, vp, vp, vp]
        l.otgpu_batch_upload_regions.argtypes = [vp, vp, u64, vp, vp, u32, vp, u32, vp, u64]
        l.otgpu_batch_upload_pairs.argtypes = [vp, vp, u64, vp, vp, u32, vp, u64]
        l.otgpu_batch_align.argtypes = [vp]
        l.otgpu_batch_download.argtypes = [vp, vp, u64]
        l.otgpu_synchronize.argtypes = [vp]
        l.otgpu_last_align_ms.argtypes = [vp, C.POINTER(C.c_float * 4)]
        l.otgpu_affine_cigars.argtypes = [vp, vp, u64, vp, vp, u32, vp, u64, i32, i32, i32, vp, vp, vp, vp]
        l.otgpu_edit_cigars.argtypes = [vp, vp, u64, vp, vp, u32, vp, u64, vp, vp, vp, vp]
        l.otgpu_affine_cigars_packed.argtypes = [vp, vp, u64, vp, vp, u32, vp, u64, i32, i32, i32, vp, vp,
                                                 C.POINTER(vp), C.POINTER(vp)]
        l.otgpu_edit_cigars_packed.argtypes = [vp, vp, u64, vp, vp, u32, vp, u64, vp, vp,
                                               C.POINTER(vp), C.POINTER(vp)]
        l.otgpu_extract_reads.argtypes = [vp, vp, u32, vp, u64, vp, u64, vp, vp, vp, C.POINTER(vp),
                                          C.POINTER(vp)]
        l.otgpu_kde_densities.argtypes = [vp, vp, vp, u32, vp, vp, u32, vp]
        l.otgpu_last_chunk_count.restype = u64
        l.otgpu_last_chunk_count.argtypes = [vp]
        l.otgpu_first_window.restype = C.c_int
        l.otgpu_first_window.argtypes = [vp]
        _lib = l
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None and a.size else None


class PinnedArray:
    """numpy view over otgpu_host_alloc memory."""

    def __init__(self, shape, dtype):
        self.dtype = np.dtype(dtype)
        n = int(np.prod(shape)) * self.dtype.itemsize
        self.ptr = C.c_void_p()
        st = lib().otgpu_host_alloc(max(n, 1), C.byref(self.ptr))
        if st != 0:
            raise OtgpuError(lib().otgpu_strerror(st).decode())
        buf = (C.c_uint8 * max(n, 1)).from_address(self.ptr.value)
        self.array = np.frombuffer(buf, dtype=self.dtype, count=int(np.prod(shape))).reshape(shape)

    def free(self):
        if self.ptr:
            self.array = None
            lib().otgpu_host_free(self.ptr)
            self.ptr = None


class Context:
    """Thin owner of an otgpu_ctx."""

    def __init__(self, device=0):
        self._l = lib()
        self._h = C.c_void_p()
        st = self._l.otgpu_create(device, C.byref(self._h))
        if st != 0:
            raise OtgpuError(self._l.otgpu_strerror(st).decode())

    def close(self):
        if self._h:
            self._l.otgpu_destroy(self._h)
            self._h = C.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def _ck(self, st):
        if st != 0:
            raise OtgpuError(f"{self._l.otgpu_strerror(st).decode()}: "
                             f"{self._l.otgpu_last_cuda_error(self._h).decode()}")

    @property
    def stream(self):
        return self._l.otgpu_stream(self._h)

    def stats(self):
        s = Stats()
        self._ck(self._l.otgpu_get_stats(self._h, C.byref(s)))
        return dict(n_pairs=s.n_pairs, n_pairs_8bit=s.n_pairs_8bit, n_pairs_wide=s.n_pairs_wide,
                    kernel_launches=s.kernel_launches)

    def edit_scores(self, bases, off, lens, pairs):
        pairs = np.ascontiguousarray(pairs, dtype=PAIR_DTYPE)
        out = np.empty(len(pairs), dtype=np.int32)
        self._ck(self._l.otgpu_edit_scores(self._h, _p(bases), bases.size, _p(off), _p(lens),
                                           len(lens), _p(pairs), len(pairs), _p(out)))
        return out

    def region_edit_scores(self, bases, off, lens, region_start, score_off, n_scores):
        out = np.empty(n_scores, dtype=np.int32)
        self._ck(self._l.otgpu_region_edit_scores(self._h, _p(bases), bases.size, _p(off), _p(lens),
                                                  len(lens), _p(region_start), len(region_start) - 1,
                                                  _p(score_off), n_scores, _p(out)))
        return out

    def region_edit_scores_hooked(self, bases, off, lens, region_start, score_off, n_scores,
                                  wait_bases=None, scores_ready=None):
        """otgpu_region_edit_scores_hooked: wait_bases(upto_byte) is called before the library first
        reads bases[:upto_byte], scores_ready(first, n) once out[first:first+n] is final."""
        WAIT = C.CFUNCTYPE(None, C.c_void_p, C.c_uint64)
        READY = C.CFUNCTYPE(None, C.c_void_p, C.c_uint64, C.c_uint64)

        class Hooks(C.Structure):
            _fields_ = [("wait_bases", WAIT), ("scores_ready", READY), ("user", C.c_void_p)]
        out = np.empty(n_scores, dtype=np.int32)
        h = Hooks(WAIT(lambda u, upto: wait_bases(int(upto)) if wait_bases else None),
                  READY(lambda u, a, n: scores_ready(int(a), int(n), out) if scores_ready else None), None)
        self._ck(self._l.otgpu_region_edit_scores_hooked(
            self._h, _p(bases), C.c_uint64(bases.size), _p(off), _p(lens), C.c_uint32(len(lens)),
            _p(region_start), C.c_uint32(len(region_start) - 1), _p(score_off), C.c_uint64(n_scores),
            _p(out), C.byref(h)))
        return out

    def region_edit_scores_extracted(self, read_index, region_start, score_off, n_scores):
        """fill_dist_matrix over records of the last extract_reads call, device to device."""
        read_index = np.ascontiguousarray(read_index, dtype=np.uint32)
        out = np.empty(n_scores, dtype=np.int32)
        self._ck(self._l.otgpu_region_edit_scores_extracted(
            self._h, _p(read_index), len(read_index), _p(region_start), len(region_start) - 1,
            _p(score_off), n_scores, _p(out)))
        return out

    def last_chunk_count(self):
        return int(self._l.otgpu_last_chunk_count(self._h))

    def first_window(self):
        """Flavour of the <= 8 kb score kernel the next call will use (otgpu_first_window)."""
        return int(self._l.otgpu_first_window(self._h))

    def edit_cigars(self, bases, off, lens, pairs):
        """Edit distance with CIGAR (compare's aligner) -> (scores, CIGAR bytes | status)."""
        return self.affine_cigars(bases, off, lens, pairs, edit=True)

    def edit_cigar_lengths(self, bases, off, lens, pairs, scores=None, clen=None):
        """Scores and CIGAR lengths only (what get_distances uses): no CIGAR bytes leave the GPU."""
        pairs = np.ascontiguousarray(pairs, dtype=PAIR_DTYPE)
        n = len(pairs)
        scores = np.zeros(n, dtype=np.int32) if scores is None else scores
        clen = np.zeros(n, dtype=np.int32) if clen is None else clen
        self._ck(self._l.otgpu_edit_cigars(self._h, _p(bases), bases.size, _p(off), _p(lens), len(lens),
                                           _p(pairs), n, _p(scores), None, None, _p(clen)))
        return scores, clen

    def kde_densities(self, values, value_off, bandwidth, grid):
        """KDE::f at every grid point for every region -> float64[n_regions, n_grid]."""
        values = np.ascontiguousarray(values, dtype=np.float64)
        value_off = np.ascontiguousarray(value_off, dtype=np.uint64)
        bandwidth = np.ascontiguousarray(bandwidth, dtype=np.float64)
        grid = np.ascontiguousarray(grid, dtype=np.float64)
        n = len(value_off) - 1
        out = np.zeros((n, len(grid)), dtype=np.float64)
        self._ck(self._l.otgpu_kde_densities(self._h, _p(values), _p(value_off), n, _p(bandwidth), _p(grid),
                                             len(grid), _p(out)))
        return out

    def extract_reads(self, records, cigar, seq4):
        """get_breakpoints + parse_alignment for a batch of BAM records (BAM_READ_DTYPE) ->
        (status uint8[n], coords int32[n,2], lens uint32[n], begin uint64[n+1], bases uint8[total]);
        the last two are views of pinned memory the context owns."""
        records = np.ascontiguousarray(records, dtype=BAM_READ_DTYPE)
        cigar = np.ascontiguousarray(cigar, dtype=np.uint32)
        seq4 = np.ascontiguousarray(seq4, dtype=np.uint8)
        n = len(records)
        status = np.zeros(n, dtype=np.uint8)
        coords = np.zeros((n, 2), dtype=np.int32)
        lens = np.zeros(n, dtype=np.uint32)
        pb, pg = C.c_void_p(), C.c_void_p()
        self._ck(self._l.otgpu_extract_reads(self._h, _p(records), n, _p(cigar), cigar.size, _p(seq4),
                                             seq4.size, _p(status), _p(coords), _p(lens),
                                             C.byref(pb), C.byref(pg)))
        if n == 0 or not pg.value:
            return status, coords, lens, np.zeros(1, np.uint64), np.zeros(0, np.uint8)
        begin = np.ctypeslib.as_array(C.cast(pg, C.POINTER(C.c_uint64)), shape=(n + 1,))
        total = int(begin[n])
        bases = np.ctypeslib.as_array(C.cast(pb, C.POINTER(C.c_uint8)), shape=(max(total, 1),))[:total]
        return status, coords, lens, begin, bases

    def cigars_packed(self, bases, off, lens, pairs, x=4, o=6, e=2, edit=False):
        """Packed output: -> (scores, lengths, begin uint64[n+1], packed uint8[total]); the last two
        are views of pinned memory the context owns, valid until the next call."""
        pairs = np.ascontiguousarray(pairs, dtype=PAIR_DTYPE)
        n = len(pairs)
        scores = np.zeros(n, dtype=np.int32)
        clen = np.zeros(n, dtype=np.int32)
        pc, pb = C.c_void_p(), C.c_void_p()
        if edit:
            self._ck(self._l.otgpu_edit_cigars_packed(self._h, _p(bases), bases.size, _p(off), _p(lens),
                                                      len(lens), _p(pairs), n, _p(scores), _p(clen),
                                                      C.byref(pc), C.byref(pb)))
        else:
            self._ck(self._l.otgpu_affine_cigars_packed(self._h, _p(bases), bases.size, _p(off), _p(lens),
                                                        len(lens), _p(pairs), n, x, o, e, _p(scores),
                                                        _p(clen), C.byref(pc), C.byref(pb)))
        if n == 0 or not pb.value:
            return scores, clen, np.zeros(1, np.uint64), np.zeros(0, np.uint8)
        begin = np.ctypeslib.as_array(C.cast(pb, C.POINTER(C.c_uint64)), shape=(n + 1,))
        total = int(begin[n])
        packed = np.ctypeslib.as_array(C.cast(pc, C.POINTER(C.c_uint8)), shape=(max(total, 1),))[:total]
        return scores, clen, begin, packed

    def affine_cigars(self, bases, off, lens, pairs, x=4, o=6, e=2, edit=False):
        """-> (scores int32[n], list of CIGAR bytes or negative status ints)."""
        pairs = np.ascontiguousarray(pairs, dtype=PAIR_DTYPE)
        n = len(pairs)
        caps = (lens[pairs["pattern"]].astype(np.uint64) + lens[pairs["text"]].astype(np.uint64)) if n \
            else np.zeros(0, np.uint64)
        coff = np.zeros(n + 1, dtype=np.uint64)
        coff[1:] = np.cumsum(caps)
        scores = np.zeros(n, dtype=np.int32)
        clen = np.zeros(n, dtype=np.int32)
        cig = np.zeros(int(coff[-1]) + 1, dtype=np.uint8)
        if edit:
            self._ck(self._l.otgpu_edit_cigars(self._h, _p(bases), bases.size, _p(off), _p(lens),
                                               len(lens), _p(pairs), n, _p(scores), _p(cig),
                                               _p(coff), _p(clen)))
        else:
            self._ck(self._l.otgpu_affine_cigars(self._h, _p(bases), bases.size, _p(off), _p(lens),
                                                 len(lens), _p(pairs), n, x, o, e, _p(scores),
                                                 _p(cig), _p(coff), _p(clen)))
        out = []
        for i in range(n):
            out.append(cig[int(coff[i]):int(coff[i]) + int(clen[i])].tobytes() if clen[i] >= 0
                       else int(clen[i]))
        return scores, out

    def upload_regions(self, bases, off, lens, region_start, score_off, n_scores):
        self._ck(self._l.otgpu_batch_upload_regions(self._h, _p(bases), bases.size, _p(off),
                                                    _p(lens), len(lens), _p(region_start),
                                                    len(region_start) - 1, _p(score_off), n_scores))

    def upload_pairs(self, bases, off, lens, pairs):
        pairs = np.ascontiguousarray(pairs, dtype=PAIR_DTYPE)
        self._ck(self._l.otgpu_batch_upload_pairs(self._h, _p(bases), bases.size, _p(off), _p(lens),
                                                  len(lens), _p(pairs), len(pairs)))

    def align(self):
        self._ck(self._l.otgpu_batch_align(self._h))

    def download(self, out):
        self._ck(self._l.otgpu_batch_download(self._h, _p(out), out.size))
        return out

    def last_align_ms(self):
        ms = (C.c_float * 4)()
        self._ck(self._l.otgpu_last_align_ms(self._h, C.byref(ms)))
        return dict(pack=ms[0], expand=ms[1], align=ms[2], wide=ms[3])

    def synchronize(self):
        self._ck(self._l.otgpu_synchronize(self._h))
