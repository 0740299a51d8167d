"""ctypes binding of libbrisk_b200.so — one prototype per entry point of include/brisk_b200.h."""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libbrisk_b200.so")

# name -> (restype, argtypes); the CPU test-suite checks that the library exports every one of them
PROTOTYPES = {
    "brisk_create": (C.c_int, [C.c_uint8, C.c_uint8, C.c_uint8, C.c_uint32, C.c_int, C.c_uint64, C.c_uint32,
                               C.POINTER(C.c_void_p)]),
    "brisk_destroy": (C.c_int, [C.c_void_p]),
    "brisk_enumerate_reads": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p, C.c_uint64,
                                        C.POINTER(C.c_uint64), C.POINTER(C.c_uint32)]),
    "brisk_enumerate_reads_tagged": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p, C.c_uint64,
                                               C.POINTER(C.c_uint64), C.POINTER(C.c_uint32), C.c_void_p]),
    "brisk_count_reads": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64]),
    "brisk_insert_reads": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p]),
    "brisk_get_reads": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p]),
    "brisk_query_reads": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]),
    "brisk_insert_kmers": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p, C.c_int]),
    "brisk_get_kmers": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p]),
    "brisk_data_ptr": (C.c_void_p, [C.c_void_p, C.c_uint64]),
    "brisk_dump": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64, C.POINTER(C.c_uint64)]),
    "brisk_dump_ordered": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64, C.POINTER(C.c_uint64)]),
    "brisk_stats": (C.c_int, [C.c_void_p] + [C.POINTER(C.c_uint64)] * 5),
    "brisk_reserve": (C.c_int, [C.c_void_p, C.c_uint64]),
    "brisk_capacity": (C.c_int, [C.c_void_p, C.POINTER(C.c_uint64)]),
    "brisk_sync": (C.c_int, [C.c_void_p]),
    "brisk_set_stream": (C.c_int, [C.c_void_p, C.c_void_p]),
    "brisk_last_timing": (C.c_int, [C.c_void_p, C.POINTER(C.c_float), C.POINTER(C.c_uint32)]),
    "brisk_parse_fasta": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p, C.c_uint64, C.c_void_p, C.c_uint64,
                                    C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]),
    "brisk_count_fasta": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint64, C.POINTER(C.c_uint64)]),
    "brisk_query_fasta": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint64, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]),
    "brisk_partition_reads": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint32, C.c_void_p, C.c_uint64,
                                        C.c_void_p, C.POINTER(C.c_uint32)]),
    "brisk_apply_records": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint64, C.c_int, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]),
    "brisk_comm_info": (C.c_int, [C.c_void_p, C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    "brisk_exchange_plan": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p]),
    "brisk_comm_unique_id": (C.c_int, [C.c_void_p]),
    "brisk_comm_init": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int]),
    "brisk_owner_of_bucket": (C.c_uint32, [C.c_uint64, C.c_uint8, C.c_uint32]),
    "brisk_write_kff": (C.c_int, [C.c_void_p, C.c_char_p, C.c_void_p, C.c_void_p]),
    "brisk_reparameterise": (C.c_int, [C.c_void_p, C.c_uint8, C.c_uint8]),
    "brisk_long_read_info": (C.c_int, [C.c_void_p, C.c_void_p]),
    "brisk_checksum": (C.c_int, [C.c_void_p, C.c_void_p]),
    "brisk_checksum_kmers_noalloc": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p]),
    "brisk_synth_genome": (C.c_int, [C.c_void_p, C.c_uint64, C.c_uint64, C.c_void_p]),
    "brisk_synth_reads": (C.c_int, [C.c_void_p, C.c_uint64, C.c_uint32, C.c_uint64, C.c_uint64, C.c_uint64, C.c_void_p, C.c_void_p]),
    "brisk_decycle_tables": (C.c_int, [C.c_uint8, C.c_uint8, C.c_uint8, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "brisk_last_error": (C.c_char_p, []),
    "brisk_version": (C.c_char_p, []),
}

_LIB = None


def load():
    """Loads the CUDA library; there is no fallback when it is missing."""
    global _LIB
    if _LIB is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                "libbrisk_b200.so is not built (run `make` or `python -c 'import __graft_entry__ as g; g.build()'`); "
                "brisk_b200 has no CPU fallback")
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in PROTOTYPES.items():
            fn = getattr(lib, name)
            fn.restype = res
            fn.argtypes = args
        _LIB = lib
    return _LIB
