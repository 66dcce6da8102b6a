// Joe-Kuo direction numbers, embedded from stock_market_monte_carlo_b200/data/sobol_joe_kuo_1111x32.bin
// (uint32 [1111][32], V[d][k] = m_{d,k} << (31-k); extracted from the reference's vendored Burkardt source by
// tools/extract_sobol_directions.py and validated there against the compiled i4_sobol).
#include <cstdint>

__asm__(
    ".section .rodata\n"
    ".balign 16\n"
    ".global smmc_sobol_table_begin\n"
    "smmc_sobol_table_begin:\n"
    ".incbin \"../data/sobol_joe_kuo_1111x32.bin\"\n"
    ".global smmc_sobol_table_end\n"
    "smmc_sobol_table_end:\n"
    ".previous\n");

extern "C" const uint32_t smmc_sobol_table_begin[];
extern "C" const uint32_t smmc_sobol_table_end[];

namespace smmc {
const uint32_t *sobol_table() { return smmc_sobol_table_begin; }
int sobol_table_dims() { return (int)((smmc_sobol_table_end - smmc_sobol_table_begin) / 32); }
}  // namespace smmc
