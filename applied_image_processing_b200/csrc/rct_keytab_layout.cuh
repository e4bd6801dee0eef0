// rct_keytab_layout.cuh -- constants and the per-slot memory block of the key table (see rct_keytab.cuh);
// separate so that k_match (rct_sort.cuh), which records the run starts, can address the block.
#pragma once

constexpr int kKtCells = 65537;                            // cell c covers the coordinate interval [c - 0.5, c + 0.5)
// The table holds 2^17 entries: the apply pass masks the rounded coordinate instead of clamping it, and every entry
// above kKtCells - 1 is flagged.  Real target colours stay below 2^16 + kKtSlackCells cells by the error bound (the
// builder refuses frames whose bound is wider); whatever an unwanted lane (background pixel, black) computes is never used.
#ifndef RCT_KT_NOCLAMP
#define RCT_KT_NOCLAMP 1
#endif
constexpr int kKtCellsAlloc = RCT_KT_NOCLAMP ? 131072 : 65537;
constexpr double kKtSlackCells = 1024.0;
constexpr int kKtRunCap = 16384;                           // run starts per frame the table is built for (else: LUT only)
constexpr unsigned kKtFlag = 0x80000000u;                  // cell is ambiguous: resolve through the colour LUT
constexpr double kKtEpsKey = 0x1p-16;
constexpr double kKtEpsC = 1.0 / 64.0;                     // error bound of the fp32 cell coordinate, in cells

struct KtParams {
    float m[9];        // RGB_TO_LMS / 255
    float w[3];        // log10(2) * (components_ @ LMS_TO_LAB)
    float pmk, tscale; // cell coordinate = (w . lg2(lms) - pmk) * tscale, pmk = projected mean + smallest target key
    unsigned valid;    // 0: the apply pass uses the colour LUT only
    unsigned n_flagged, n_runs;
    unsigned long long kmax_enc, kmin_enc_inv;  // ord_enc(largest target key), ~ord_enc(smallest): atomic max by k_match
    double kmin, scale;                         // the builder's exact coordinate: (key - kmin) * scale
    unsigned pad_[38];
};
static_assert(sizeof(KtParams) == 256, "KtParams is one 256-byte block");

// per-slot block: KtParams | seeds uint64[kKtCells] ((sorted index + 1) << 32 | colour of the run whose widened threshold
// range ends just before the cell) | flags uint8[kKtCells] | cells uint32[kKtCells] (colour, or kKtFlag)
// | run keys double[kKtRunCap] | run (sorted index, colour) uint2[kKtRunCap].  KtParams, seeds and flags are cleared before
// every build (kKtZeroBytes from the start of the block).
constexpr size_t kt_align(size_t v) { return (v + 255) / 256 * 256; }
constexpr size_t kKtOffSeeds = 256;
constexpr size_t kKtOffFlags = kKtOffSeeds + kt_align((size_t)kKtCells * 8);
constexpr size_t kKtOffCells = kKtOffFlags + kt_align((size_t)kKtCells);
constexpr size_t kKtOffRunKey = kKtOffCells + kt_align((size_t)kKtCellsAlloc * 4);
constexpr size_t kKtOffRunIc = kKtOffRunKey + (size_t)kKtRunCap * 8;
constexpr size_t kKtBytes = kKtOffRunIc + (size_t)kKtRunCap * 8;
constexpr size_t kKtZeroBytes = kKtOffCells;
