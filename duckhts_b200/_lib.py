"""ctypes binding of libduckhts_cuda.so (include/duckhts_cuda.h).  Fails loudly if the library is absent."""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
lib_path = os.path.join(_HERE, "lib", "libduckhts_cuda.so")


class DuckhtsError(RuntimeError):
    pass


class BgzfBlock(C.Structure):
    _fields_ = [("coffset", C.c_uint64), ("csize", C.c_uint32), ("isize", C.c_uint32), ("crc", C.c_uint32),
                ("reserved", C.c_uint32)]


class InflateStats(C.Structure):
    _fields_ = [("n_blocks", C.c_uint64), ("comp_bytes", C.c_uint64), ("out_bytes", C.c_uint64),
                ("bad_block", C.c_int64), ("bad_code", C.c_int32), ("eof_marker", C.c_int32),
                ("kernel_ms", C.c_float), ("total_ms", C.c_float)]


class DeflateStats(C.Structure):
    _fields_ = [("in_bytes", C.c_uint64), ("out_bytes", C.c_uint64), ("n_blocks", C.c_uint64), ("h2d_bytes", C.c_uint64),
                ("d2h_bytes", C.c_uint64), ("kernel_ms", C.c_float), ("total_ms", C.c_float)]


class TextBatch(C.Structure):
    _fields_ = [("n_lines", C.c_uint64), ("line0", C.c_uint64), ("text", C.c_void_p), ("text_bytes", C.c_uint64),
                ("cells", C.c_void_p), ("flags", C.c_void_p), ("n_proj", C.c_uint32), ("status", C.c_int32)]


DH_BAM_CORE_COLS = 13
DH_BAM_MAX_COLS = 13 + 63


class TagReq(C.Structure):
    _fields_ = [("tag", C.c_char * 2), ("type", C.c_char), ("pad", C.c_char)]


class BamBatch(C.Structure):
    _fields_ = [("n_rows", C.c_uint64), ("row0", C.c_uint64), ("status", C.c_int32), ("n_cols", C.c_int32),
                ("data", C.c_void_p * DH_BAM_MAX_COLS), ("validity", C.c_void_p * DH_BAM_MAX_COLS),
                ("heap", C.c_void_p), ("heap_bytes", C.c_uint64), ("bytes", C.c_uint64),
                ("aux_map", C.c_void_p), ("aux_map_validity", C.c_void_p), ("aux_keys", C.c_void_p), ("aux_values", C.c_void_p),
                ("n_aux", C.c_uint64), ("slot", C.c_int32), ("layout", C.c_uint32), ("rec_off", C.c_void_p)]


class BamScanStats(C.Structure):
    _fields_ = [("rows", C.c_uint64), ("h2d_bytes", C.c_uint64), ("d2h_bytes", C.c_uint64), ("inflated_bytes", C.c_uint64),
                ("n_blocks", C.c_uint64), ("rewalk_rounds", C.c_uint64), ("status", C.c_int32), ("pad", C.c_int32),
                ("end_voffset", C.c_uint64), ("inflate_ms", C.c_float), ("walk_ms", C.c_float), ("decode_ms", C.c_float)]


class BcfField(C.Structure):
    _fields_ = [("name", C.c_char_p), ("type", C.c_int32), ("is_list", C.c_int32)]


class BcfBatch(C.Structure):
    _fields_ = [("n_records", C.c_uint64), ("n_samples", C.c_uint64), ("rec0", C.c_uint64), ("status", C.c_int32), ("pad", C.c_int32),
                ("core", C.c_void_p * 7), ("core_validity", C.c_void_p * 7),
                ("info", C.POINTER(C.c_void_p)), ("info_validity", C.POINTER(C.c_void_p)),
                ("format", C.POINTER(C.c_void_p)), ("format_validity", C.POINTER(C.c_void_p)),
                ("heap", C.c_void_p), ("heap_bytes", C.c_uint64), ("bytes", C.c_uint64), ("slot", C.c_int32), ("layout", C.c_uint32),
                ("format_rec", C.POINTER(C.c_void_p))]


_lib = None

# name -> (restype, argtypes); tests/test_abi.py checks every symbol declared in include/*.h is here and exported
SIGNATURES = {
    "dh_version": (C.c_char_p, []),
    "dh_last_error": (C.c_char_p, []),
    "dh_device_count": (C.c_int, []),
    "dh_trim": (None, []),
    "dh_ctx_create": (C.c_void_p, [C.c_int]),
    "dh_ctx_destroy": (None, [C.c_void_p]),
    "dh_ctx_sm_count": (C.c_int, [C.c_void_p]),
    "dh_host_alloc": (C.c_void_p, [C.c_size_t]),
    "dh_host_free": (None, [C.c_void_p]),
    "dh_bgzf_walk": (C.c_int64, [C.c_void_p, C.c_uint64, C.c_uint64, C.POINTER(BgzfBlock), C.c_int64,
                                 C.POINTER(C.c_uint64), C.c_int]),
    "dh_bgzf_inflate_host": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p, C.c_uint64,
                                       C.POINTER(C.c_uint64), C.POINTER(InflateStats)]),
    "dh_bgzf_dev_load": (C.c_void_p, [C.c_void_p, C.c_void_p, C.c_uint64]),
    "dh_bgzf_dev_free": (None, [C.c_void_p]),
    "dh_bgzf_dev_nblocks": (C.c_uint64, [C.c_void_p]),
    "dh_bgzf_dev_comp_bytes": (C.c_uint64, [C.c_void_p]),
    "dh_bgzf_dev_out_bytes": (C.c_uint64, [C.c_void_p]),
    "dh_bgzf_dev_inflate": (C.c_int, [C.c_void_p]),
    "dh_bgzf_dev_check": (C.c_int, [C.c_void_p, C.POINTER(C.c_int64), C.POINTER(C.c_int32)]),
    "dh_bgzf_dev_read": (C.c_int, [C.c_void_p, C.c_uint64, C.c_void_p, C.c_uint64]),
    "dh_bam_open": (C.c_void_p, [C.c_void_p, C.c_void_p, C.c_uint64]),
    "dh_bam_open_file": (C.c_void_p, [C.c_void_p, C.c_char_p]),
    "dh_bam_close": (None, [C.c_void_p]),
    "dh_bam_n_ref": (C.c_int32, [C.c_void_p]),
    "dh_bam_ref_name": (C.c_char_p, [C.c_void_p, C.c_int32]),
    "dh_bam_ref_len": (C.c_uint32, [C.c_void_p, C.c_int32]),
    "dh_bam_header_text": (C.c_void_p, [C.c_void_p, C.POINTER(C.c_uint32)]),
    "dh_bam_first_voffset": (C.c_uint64, [C.c_void_p]),
    "dh_bam_n_blocks": (C.c_uint64, [C.c_void_p]),
    "dh_bam_set_projection": (C.c_int, [C.c_void_p, C.c_uint64, C.POINTER(TagReq), C.c_int]),
    "dh_bam_set_aux_map": (C.c_int, [C.c_void_p, C.c_int, C.c_char_p, C.c_int]),
    "dh_bam_configure": (C.c_int, [C.c_void_p, C.c_uint64, C.c_uint32, C.c_uint64]),
    "dh_bam_set_mode": (C.c_int, [C.c_void_p, C.c_uint32]),
    "dh_bam_scan_begin": (C.c_int, [C.c_void_p, C.c_uint64, C.c_uint64]),
    "dh_bam_next_batch": (C.c_int, [C.c_void_p, C.POINTER(BamBatch)]),
    "dh_bam_set_slots": (C.c_int, [C.c_void_p, C.c_int]),
    "dh_bam_prefetch": (C.c_int, [C.c_void_p]),
    "dh_bam_acquire_batch": (C.c_int, [C.c_void_p, C.POINTER(BamBatch)]),
    "dh_bam_release_batch": (C.c_int, [C.c_void_p, C.c_int32]),
    "dh_bam_batch_checksum": (C.c_int, [C.c_void_p, C.POINTER(BamBatch), C.POINTER(C.c_uint64)]),
    "dh_bam_n_rg": (C.c_int32, [C.c_void_p]),
    "dh_bam_rg_id": (C.c_void_p, [C.c_void_p, C.c_int32, C.POINTER(C.c_uint32)]),
    "dh_bam_rg_sample": (C.c_char_p, [C.c_void_p, C.c_int32]),
    "dh_bam_clone": (C.c_void_p, [C.c_void_p, C.c_void_p]),
    "dh_bam_set_filter": (C.c_int, [C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32]),
    "dh_bam_index_build": (C.c_int, [C.c_void_p, C.POINTER(C.c_void_p), C.POINTER(C.c_uint64)]),
    "dh_free": (None, [C.c_void_p]),
    "dh_index_split_contigs": (C.c_int, [C.c_void_p, C.c_int, C.POINTER(C.c_int32)]),
    "dh_bgzf_inflate_fd": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.POINTER(InflateStats)]),
    "dh_bgzf_deflate_bound": (C.c_uint64, [C.c_uint64]),
    "dh_text_open_file": (C.c_void_p, [C.c_void_p, C.c_char_p]),
    "dh_text_close": (None, [C.c_void_p]),
    "dh_text_head": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint64, C.POINTER(C.c_uint64)]),
    "dh_text_configure": (C.c_int, [C.c_void_p, C.c_uint32, C.POINTER(C.c_uint32), C.c_uint32, C.c_int]),
    "dh_text_scan_begin": (C.c_int, [C.c_void_p]),
    "dh_text_next_batch": (C.c_int, [C.c_void_p, C.POINTER(TextBatch)]),
    "dh_bgzf_deflate_host": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint64, C.c_int, C.c_void_p, C.c_uint64, C.POINTER(C.c_uint64),
                                       C.POINTER(DeflateStats)]),
    "dh_bgzf_deflate_fd": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.POINTER(DeflateStats)]),
    "dh_bcf_set_slots": (C.c_int, [C.c_void_p, C.c_int]),
    "dh_bcf_acquire_batch": (C.c_int, [C.c_void_p, C.POINTER(BcfBatch)]),
    "dh_bcf_release_batch": (C.c_int, [C.c_void_p, C.c_int32]),
    "dh_bam_scan_stats_get": (C.c_int, [C.c_void_p, C.POINTER(BamScanStats)]),
    "dh_bam_split": (C.c_int, [C.c_void_p, C.c_int, C.POINTER(C.c_uint64)]),
    "dh_index_load": (C.c_void_p, [C.c_void_p, C.c_void_p, C.c_uint64]),
    "dh_index_free": (None, [C.c_void_p]),
    "dh_index_n_ref": (C.c_int32, [C.c_void_p]),
    "dh_index_is_csi": (C.c_int, [C.c_void_p]),
    "dh_index_n_no_coor": (C.c_uint64, [C.c_void_p]),
    "dh_bam_scan_regions": (C.c_int, [C.c_void_p, C.c_void_p, C.POINTER(C.c_char_p), C.c_int]),
    "dh_bam_scan_contigs": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int32, C.c_int32]),
    "dh_bam_split_contigs": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.POINTER(C.c_int32)]),
    "dh_bcf_open": (C.c_void_p, [C.c_void_p, C.c_void_p, C.c_uint64]),
    "dh_bcf_close": (None, [C.c_void_p]),
    "dh_bcf_n_samples": (C.c_int32, [C.c_void_p]),
    "dh_bcf_sample_name": (C.c_char_p, [C.c_void_p, C.c_int32]),
    "dh_bcf_n_contigs": (C.c_int32, [C.c_void_p]),
    "dh_bcf_contig_name": (C.c_char_p, [C.c_void_p, C.c_int32]),
    "dh_bcf_n_info": (C.c_int32, [C.c_void_p]),
    "dh_bcf_n_format": (C.c_int32, [C.c_void_p]),
    "dh_bcf_info_field": (C.c_int, [C.c_void_p, C.c_int32, C.POINTER(BcfField)]),
    "dh_bcf_format_field": (C.c_int, [C.c_void_p, C.c_int32, C.POINTER(BcfField)]),
    "dh_bcf_has_vep": (C.c_int, [C.c_void_p]),
    "dh_bcf_header_text": (C.c_void_p, [C.c_void_p, C.POINTER(C.c_uint32)]),
    "dh_bcf_set_projection": (C.c_int, [C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p]),
    "dh_bcf_configure": (C.c_int, [C.c_void_p, C.c_uint64, C.c_uint32, C.c_uint64]),
    "dh_bcf_set_mode": (C.c_int, [C.c_void_p, C.c_uint32]),
    "dh_bcf_scan_begin": (C.c_int, [C.c_void_p]),
    "dh_bcf_scan_region": (C.c_int, [C.c_void_p, C.c_void_p, C.c_char_p]),
    "dh_bcf_next_batch": (C.c_int, [C.c_void_p, C.POINTER(BcfBatch)]),
    "dh_bcf_scan_stats_get": (C.c_int, [C.c_void_p, C.POINTER(BamScanStats)]),
    "dh_ctx_sync": (C.c_int, [C.c_void_p]),
    "dh_timer_start": (C.c_int, [C.c_void_p]),
    "dh_timer_stop": (C.c_int, [C.c_void_p, C.POINTER(C.c_float)]),
    "dh_flush_l2": (C.c_int, [C.c_void_p]),
    "dh_measure_pcie": (C.c_int, [C.c_void_p, C.c_uint64, C.POINTER(C.c_float), C.POINTER(C.c_float)]),
    "dh_bgzf_dev_inflate_timed": (C.c_int, [C.c_void_p, C.POINTER(C.c_float)]),
    "dh_measure_pcie_bidir": (C.c_int, [C.c_void_p, C.c_uint64, C.POINTER(C.c_float), C.POINTER(C.c_float)]),
    "dh_ctx_launch_count": (C.c_uint64, [C.c_void_p]),
    "dh_bgzf_inflate_noalloc": (C.c_int, [C.c_int, C.c_void_p, C.c_uint64, C.POINTER(BgzfBlock), C.c_uint32, C.c_void_p, C.c_uint64,
                                          C.POINTER(C.c_uint64), C.POINTER(C.c_int64), C.POINTER(C.c_int32), C.c_void_p]),
}


def lib():
    """Load libduckhts_cuda.so once.  No fallback: a missing library is an error, not a slow path."""
    global _lib
    if _lib is None:
        if not os.path.exists(lib_path):
            raise DuckhtsError(f"{lib_path} not built: run `python -c 'import __graft_entry__ as g; g.build()'` "
                               "(duckhts_b200 has no CPU fallback)")
        _lib = C.CDLL(lib_path)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(_lib, name)
            fn.restype = res
            fn.argtypes = args
    return _lib


def check(rc, what=""):
    if rc is None or (isinstance(rc, int) and rc < 0):
        msg = lib().dh_last_error().decode(errors="replace")
        raise DuckhtsError(f"{what}: rc={rc}: {msg}")
    return rc
