"""Calls every entry point of include/pose_b200.h with null pointers and zero sizes (run as `python -m tests.nullfuzz_cabi`,
in a process of its own: a crash must not take the test runner with it).  The C ABI's contract (SURVEY 8b "Errors"): status
codes, never a fault.  Prints one line per call and `nullfuzz: N calls, all returned` at the end."""
import ctypes as C
import importlib
import sys

SKIP = {"pe_last_error", "pe_version", "pe_launch_count",                       # no status to check
        "pe_net_destroy", "pe_db_destroy", "pe_comm_destroy", "pe_mesh_destroy", "pe_miner_destroy", "pe_host_free",
        "pe_timer_destroy"}                                                     # destroy(null) is covered below


# zero work is a valid request there (no views, no labels, free(null), a 0-byte copy)
ZERO_WORK_OK = {"pe_template_samples", "pe_pose_labels", "pe_dev_free", "pe_memcpy_h2d", "pe_memcpy_d2h", "pe_dev_memset"}


def main():
    capi = importlib.import_module("3dposeestimation_b200.capi")
    lib = capi.lib()
    n_calls = 0
    accepted = []
    for name, (res, args) in capi.SIGNATURES.items():
        if name in SKIP:
            continue
        vals = []
        for a in args:
            if a in (C.c_float, C.c_double):
                vals.append(a(0.0))
            elif a in (C.c_int, C.c_int64, C.c_uint64, C.c_size_t, C.c_uint32):
                vals.append(a(0))
            else:
                vals.append(None)
        print(name, end=" ", flush=True)
        r = getattr(lib, name)(*vals)
        print("->", r, flush=True)
        n_calls += 1
        # 0 only where zero work is a valid request
        if r == 0 and name not in ZERO_WORK_OK:
            accepted.append(name)
    for name in sorted(SKIP):
        if name.endswith("_destroy") or name == "pe_host_free":
            print(name, "(null)", flush=True)
            getattr(lib, name)(None)
            n_calls += 1
    if accepted:
        print("nullfuzz: accepted null arguments:", " ".join(accepted))
        return 1
    print(f"nullfuzz: {n_calls} calls, all returned")
    return 0


if __name__ == "__main__":
    sys.exit(main())
