// hostcheck_selftest.cpp — drives the launcher's host reference loops (hostcheck.cpp) over ragged sizes
// on exact-size heap blocks and prints one checksum line per case.  Built plain and with
// -fsanitize=address,undefined (`make hostcheck-selftest`); tests/test_launcher_cpu.py requires both
// outputs to agree with each other and with the oracle.  Not part of the pod image.
#include <cinttypes>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "hostcheck.h"

using namespace hostcheck;

int main()
{
    const size_t sizes[] = {0, 1, 2, 7, 4099, 100003};
    const float s = -0.333251953125f;
    const char* opn[] = {"add", "copy", "scale", "triad"};
    for (int dtype : {F32, BF16}) {
        const size_t esz = dtype == F32 ? 4 : 2;
        for (int mode = 0; mode < 3; ++mode) {
            for (size_t n : sizes) {
                unsigned char* x = static_cast<unsigned char*>(malloc(n * esz + 1));
                unsigned char* y = static_cast<unsigned char*>(malloc(n * esz + 1));
                unsigned char* d = static_cast<unsigned char*>(malloc(n * esz + 1));
                if (!x || !y || !d) return 2;
                const uint64_t start = dtype == F32 ? (uint64_t(1) << 32) - 5 : 11;
                fill(dtype, x, start, n, 7, 0, mode);
                fill(dtype, y, start, n, 7, 1, mode);
                for (int op = 0; op < 4; ++op) {
                    for (bool par : {false, true}) {
                        apply(op, dtype, d, x, y, s, n, par);
                        uint64_t c[2];
                        checksum(d, n * esz, 3, c);
                        printf("%s %s mode=%d n=%zu %016" PRIx64 ":%016" PRIx64 "\n", opn[op], dtype == F32 ? "f32" : "bf16", mode, n, c[0],
                               c[1]);
                    }
                    // the chunked "expected" path must agree with the materialised one (word offset 0, even start for bf16)
                    uint64_t e[2], m[2];
                    const uint64_t st = 12;
                    expected_checksum(op, dtype, st, n, 7, mode, s, 0, e);
                    std::vector<unsigned char> gx(n * esz + 1), gy(n * esz + 1), gd(n * esz + 1);
                    fill(dtype, gx.data(), st, n, 7, 0, mode);
                    fill(dtype, gy.data(), st, n, 7, 1, mode);
                    apply(op, dtype, gd.data(), gx.data(), gy.data(), s, n, false);
                    checksum(gd.data(), n * esz, 0, m);
                    if (e[0] != m[0] || e[1] != m[1]) {
                        printf("MISMATCH expected_checksum %s n=%zu\n", opn[op], n);
                        return 1;
                    }
                }
                free(x); free(y); free(d);
            }
        }
    }
    printf("hostcheck selftest ok\n");
    return 0;
}
