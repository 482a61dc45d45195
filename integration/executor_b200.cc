// executor_b200.cc -- the reference-side binding of INTEGRATION.md path A.
//
// Drop this file into the reference tree IN PLACE OF executor.cc (makefile: replace executor.o by executor_b200.o, add
// -I$(GSPARSE)/include -L$(GSPARSE)/generalsparse_b200/lib -lgsparse_b200 -Wl,-rpath,$(GSPARSE)/generalsparse_b200/lib).  Nothing else
// of the reference changes: its operators build the metadata, its code_generator::generate_final_program writes data_source/<id>/, its
// drivers call execute_binary(path, matrix, time, gflops) (executor.hpp:15) -- and that call now runs the B200 kernels on the files the
// reference wrote instead of compiling and forking the generated program (executor.cc:6-105).
//
// oracle/Makefile builds exactly this against the unmodified reference objects (oracle/_ref/ref_path_a); tests/test_path_a.py runs it.
#include "executor.hpp"      // the reference's own header: the two signatures below are declared there
#include "gsparse_b200.h"

#include <cstdio>
#include <cstdlib>
#include <cstring>

bool compile_spmv_code(string) { return true; } // executor.cc:60-105 ran make_kernel.sh here: nothing to compile any more

bool execute_binary(string execute_path, string filename, float &total_execute_time, float &gflops)
{
    // GS_PATH_A_KERNEL: auto (default: the library picks the B200 kernel for the chain), chain (the chain's own reference-shaped
    // family), or a kernel name; GS_PATH_A_REPEAT: launches to time (the reference's generated program bakes in 10000)
    int kernel = GS_KERNEL_AUTO;
    if (const char *k = getenv("GS_PATH_A_KERNEL")) {
        if (!strcmp(k, "chain")) kernel = GS_KERNEL_CHAIN;
        else if (strcmp(k, "auto"))
            for (int f = 0; f < GS_FAMILY_COUNT; f++)
                if (!strcmp(k, gs_family_name(f))) kernel = f;
    }
    const char *r = getenv("GS_PATH_A_REPEAT");
    const int repeat = r && atoi(r) > 0 ? atoi(r) : 10000;
    total_execute_time = 0, gflops = 0;
    if (gs_execute_binary(execute_path.c_str(), filename.c_str(), repeat, kernel, &total_execute_time, &gflops) != GS_OK) {
        fprintf(stderr, "execute_binary: %s\n", gs_last_error());
        return false; // the reference returns false on a failed run too (executor.cc:16-19,31-35)
    }
    return true;
}
