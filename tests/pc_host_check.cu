// Test harness (CPU only, built by tests/test_pc_host.py): the z-marched interpolation state machine of the fused
// prolongation + post-smoothing pass (gs_device.cuh PcInterp, used by kernels_fused.cu smooth2u_tma<.., PC>) against the
// closed-form prolong_point of kernels_simple.cu (itself bit-identical to CpuSolver::interpolate, tested elsewhere), for
// every interior point of a grid, from every chunk start parity.  Nothing here is linked into the library.
#include "../gpu-solve_b200/csrc/kernels_simple.cu"

#include <math.h>
#include <string.h>

#include <vector>

static Geom make_geom(int nx, int ny, int nz)
{
    Geom g;
    g.nx = nx; g.ny = ny; g.nz = nz;
    g.X = nx + 2; g.Y = ny + 2; g.Zg = nz + 2;
    g.z0 = 0; g.nzl = g.Zg;
    g.pitch = ((GS_XOFF + g.X + 15) / 16) * 16;
    g.plane = g.pitch * g.Y;
    return g;
}

// returns the number of mismatching (or unchecked-but-expected) values; *checked = comparisons made
extern "C" long pc_host_check(int nx, int ny, int nz, long* checked)
{
    const Geom gf = make_geom(nx, ny, nz), gc = make_geom(nx / 2, ny / 2, nz / 2);
    std::vector<double> cv((size_t)gc.plane * gc.nzl, 0.0);
    unsigned long long s = 88172645463325252ull;
    for (int z = 1; z <= gc.nz; z++)
        for (int y = 1; y <= gc.ny; y++)
            for (int x = 1; x <= gc.nx; x++) {
                s ^= s << 13; s ^= s >> 7; s ^= s << 17;
                cv[gs_idx(gc, x, y, z)] = (double)(s >> 11) / 9007199254740992.0 - 0.5;
            }
    long bad = 0, n = 0;
    for (int zstart = -2; zstart <= 3; zstart++)          // global plane of q = 0 of a chunk (both parities, below the array too)
        for (int x = 1; x <= nx; x++)
            for (int ya = 0; ya <= ny; ya += 2) {            // the even row of a row pair (ya = 0: the halo row, pair (0, 1))
                const bool pxodd = x & 1;
                const int pco = GS_XOFF + std::min(std::max(x >> 1, 0), gc.X - 2) + (int)gc.pitch * std::min(std::max(ya >> 1, 0), gc.Y - 2);
                auto ld = [&](int k, int row) {
                    const double* c = cv.data() + (i64)std::min(std::max(k - gc.z0, 0), gc.nzl - 1) * gc.plane + pco + row * (int)gc.pitch;
                    return pxodd ? 0.5 * c[0] + 0.5 * c[1] : c[0];
                };
                const bool want_m = ya - 1 >= 1;
                PcInterp pci{0.0, 0.0};
                pci.init(zstart, ld);
                for (int zg = zstart; zg <= nz + 2; zg++) {
                    double eA, eB, eJ, eM;
                    pci.step(zg, want_m, ld, eA, eB, eJ, eM);
                    if (zg < 1 || zg > nz) continue;
                    const struct { int y; double e; bool on; } rows[4] = {{ya, eA, ya >= 1}, {ya + 1, eB, ya + 1 <= ny}, {ya + 2, eJ, ya + 2 <= ny},
                                                                          {ya - 1, eM, want_m}};
                    for (const auto& r : rows) {
                        if (!r.on) continue;
                        const double want = prolong_point(cv.data(), gc, x, r.y, zg);
                        n++;
                        if (memcmp(&want, &r.e, 8) != 0 && !(want == 0.0 && r.e == 0.0)) bad++;
                    }
                }
            }
    *checked = n;
    return bad;
}
