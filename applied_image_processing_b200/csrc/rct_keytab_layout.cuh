// rct_keytab_layout.cuh -- constants and the per-slot memory block of the key table (see rct_keytab.cuh);
// separate so that k_match (rct_sort.cuh), which records the run starts, can address the block.
#pragma once

constexpr int kKtCoarse = kCoarse + 1;            // coarse cells, round-to-nearest convention: t in [cb-0.5, cb+0.5)
constexpr int kKtSubMax = 256;                    // sub-cells per coarse cell (cell width >= 2 * kKtEpsT)
constexpr int kKtCellsTarget = 65536;
constexpr int kKtCellsMax = kKtCellsTarget + kKtCoarse;
constexpr int kKtRunCap = 8192;                   // run starts per frame the table is built for (else: LUT only)
constexpr unsigned kKtFlag = 0x80000000u;         // cell is ambiguous: resolve through the colour LUT
constexpr double kKtEpsKey = 0x1p-16;
constexpr double kKtEpsT = 0x1p-9;

struct KtParams {
    float m[9];        // RGB_TO_LMS / 255
    float w[3];        // log10(2) * (components_ @ LMS_TO_LAB)
    float pm;          // mean_ @ components_.T
    float tscale, toff;
    unsigned valid;    // 0: the apply pass uses the colour LUT only
    unsigned n_cells, n_flagged, n_runs, n_sorted;
    unsigned pad_[44];
};
static_assert(sizeof(KtParams) == 256, "KtParams is one 256-byte block");

// per-slot block: KtParams | cell directory uint2[kKtCoarse] (x = first cell, y = float bits of the sub-cell count)
//               | cells uint32[kKtCellsMax] | run keys double[kKtRunCap] | run colours uint32[kKtRunCap]
constexpr size_t kKtOffDir = 256;
constexpr size_t kKtOffCells = kKtOffDir + (((size_t)kKtCoarse * 8 + 255) / 256) * 256;
constexpr size_t kKtOffRunKey = kKtOffCells + (((size_t)kKtCellsMax * 4 + 255) / 256) * 256;
constexpr size_t kKtOffRunCol = kKtOffRunKey + (size_t)kKtRunCap * 8;
constexpr size_t kKtBytes = kKtOffRunCol + (size_t)kKtRunCap * 4;

