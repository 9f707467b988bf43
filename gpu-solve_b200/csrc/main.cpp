// GpuSolve-cuda <config> — the reference driver's contract (src/main.cpp:15-114)
// on top of the C ABI: same usage line, same config format, same stdout lines, so
// it drops into runExperiments.py's `impls` list next to GpuSolve-cpu / -gtx / -sycl.
#include <sys/stat.h>

#include <iostream>

#include "../../include/gpusolve_cuda.h"

int main(int argc, char* argv[])
{
    if (argc < 2) {
        std::cerr << "Missing config file. Usage program.exe path/to/config.conf\n";
        return 1;
    }
    const char* path = argv[1];
    struct stat st;
    if (stat(path, &st) != 0 || !S_ISREG(st.st_mode)) {
        std::cerr << '"' << path << '"' << " does not exist or is not a file\n";
        return 1;
    }
    std::cout << "Using config file " << '"' << path << '"' << '\n';

    gs_params p;
    if (gs_parse_config(path, &p) != GS_OK) {
        std::cerr << gs_last_error() << '\n';
        return 1;
    }
    if (p.mode == GS_MODE_LINEAR) std::cout << "Solving linear problem\n";
    else if (p.mode == GS_MODE_NONLINEAR) std::cout << "Solving nonlinear problem\n";
    else std::cout << "Solving newton problem\n";

    int device = 0;
    if (const char* e = getenv("GPUSOLVE_DEVICE")) device = atoi(e);
    gs_handle* h = nullptr;
    // the SYCL path of the reference prints the exception and still exits 0 (src/main.cpp:107-113)
    if (gs_create(&p, device, &h) != GS_OK) {
        std::cerr << "Exception: " << gs_last_error() << '\n';
        return 0;
    }
    if (gs_solve(h, nullptr) != GS_OK) std::cerr << "Exception: " << gs_last_error() << '\n';
    if (const char* dump = getenv("GPUSOLVE_DUMP")) {
        int field = p.mode == GS_MODE_NEWTON ? GS_FIELD_NEWTONV : GS_FIELD_V;
        if (gs_dump_field(h, 0, field, dump) != GS_OK) std::cerr << gs_last_error() << '\n';
    }
    gs_destroy(h);
    return 0;
}
