// Stand-alone check of csrc/rct_jpeg.cuh.
//   jpeg_selftest --host H W in.rgb out.jpg [quality]   the __host__ __device__ bodies run serially on the CPU
//   jpeg_selftest --gpu  H W in.rgb out.jpg [quality]   the kernels on device 0 (no torch: starts in a second)
// Test infrastructure: tests/test_jpeg_device_core.py compares the output with oracle/jpeg_oracle.py and Pillow.
#include <cstdio>
#include <cstdlib>
#include <algorithm>
#include <string>
#include <vector>
#include <cuda_runtime.h>
#include "../applied_image_processing_b200/csrc/rct_jpeg.cuh"
#include "../applied_image_processing_b200/csrc/rct_jpeg_dec.cuh"

// the kernels' __host__ __device__ bodies run serially on the host (test infrastructure, never a product path)
namespace jpeg {
inline size_t encode_host_emulation(const uint8_t *rgb, int H, int W, int quality, uint8_t *out, size_t out_cap)
{
    const Geom g = make_geom(H, W, 0);
    Tables tab;
    Header hdr;
    make_tables(quality, &tab, &hdr, H, W);
    const Workspace lay = carve(g, 1, nullptr);
    short *coef = new short[(size_t)g.nblk * 64u];
    unsigned long long *offs = new unsigned long long[g.nblk];
    unsigned *stream = new unsigned[lay.stream_words]();
    for (unsigned i = 0; i < g.nblk; ++i) {
        const unsigned mcu = i / 6u;
        const int b = (int)(i % 6u);
        short *o = coef + (size_t)i * 64u;
        if (is_dummy(g, mcu, b)) {
            memset(o, 0, 128);
        } else {
            int blk[64];
            block_samples(rgb, g, mcu, b, blk);
            fdct_quantise(blk, tab.q8[b < 4 ? 0 : 1], tab.rq[b < 4 ? 0 : 1], o);
        }
    }
    // as k_jpeg_dct / k_jpeg_bits split it: AC bits + DC per block first, then the DC differences
    unsigned *meta = new unsigned[g.nblk];
    for (unsigned i = 0; i < g.nblk; ++i) {
        const int b = (int)(i % 6u), c = b < 4 ? 0 : 1;
        const short *o = coef + (size_t)i * 64u;
        meta[i] = (ac_code_bits(o, is_dummy(g, i / 6u, b), tab.dc[c], tab.ac[c]) << 16) | (unsigned)(unsigned short)o[0];
    }
    unsigned long long run = 0;
    for (unsigned i = 0; i < g.nblk; ++i) {
        const unsigned mcu = i / 6u;
        const int b = (int)(i % 6u), c = b < 4 ? 0 : 1;
        CountSink sink;
        encode_block(sink, coef + (size_t)i * 64u, is_dummy(g, mcu, b), resolved_dc(coef, g, mcu, b) - dc_predictor(coef, g, mcu, b),
                     tab.dc[c], tab.ac[c]);
        const unsigned fast = (meta[i] >> 16) + dc_code_bits(resolved_dc_meta(meta, g, mcu, b) - dc_predictor_meta(meta, g, mcu, b), tab.dc[c]);
        if (fast != sink.bits) {
            fprintf(stderr, "bit count of block %u: %u from (AC bits, DC), %u from the coefficients\n", i, fast, sink.bits);
            exit(5);
        }
        offs[i] = run;
        run += fast;
    }
    delete[] meta;
    for (unsigned k = g.nblk; k-- > 0;) {  // any order: the blocks only OR into the stream
        const unsigned mcu = k / 6u;
        const int b = (int)(k % 6u), c = b < 4 ? 0 : 1;
        StreamSink sink(stream, offs[k]);
        encode_block(sink, coef + (size_t)k * 64u, is_dummy(g, mcu, b), resolved_dc(coef, g, mcu, b) - dc_predictor(coef, g, mcu, b),
                     tab.dc[c], tab.ac[c]);
        if (k == g.nblk - 1u) {
            const int pad = (int)((8ull - (sink.position() & 7ull)) & 7ull);
            if (pad) sink.put((1u << pad) - 1u, pad);
        }
        sink.finish();
    }
    const size_t nbytes = (size_t)((run + 7ull) / 8ull);
    const unsigned char *sb = reinterpret_cast<const unsigned char *>(stream);
    size_t pos = 0;
    auto put = [&](unsigned char v) {
        if (pos < out_cap) out[pos] = v;
        ++pos;
    };
    for (int k = 0; k < hdr.len; ++k) put(hdr.b[k]);
    for (size_t k = 0; k < nbytes; ++k) {
        put(sb[k]);
        if (sb[k] == 0xFF) put(0);
    }
    put(0xFF), put(0xD9);
    delete[] coef;
    delete[] offs;
    delete[] stream;
    return pos;
}
}  // namespace jpeg

#define CK(x)                                                                                   \
    do {                                                                                        \
        cudaError_t e_ = (x);                                                                   \
        if (e_ != cudaSuccess) {                                                                \
            fprintf(stderr, "%s: %s\n", #x, cudaGetErrorString(e_));                            \
            return 2;                                                                           \
        }                                                                                       \
    } while (0)

int main(int argc, char **argv)
{
    if (argc >= 4 && (std::string(argv[1]) == "--host-decode" || std::string(argv[1]) == "--gpu-decode")) {
        // jpeg_selftest --host-decode|--gpu-decode in.jpg out.raw [par]   (prints "H W channels"; par = self-synchronising decode)
        FILE *fi = fopen(argv[2], "rb");
        if (!fi) return 1;
        std::vector<unsigned char> file;
        unsigned char buf[65536];
        for (size_t got; (got = fread(buf, 1, sizeof buf, fi)) > 0;) file.insert(file.end(), buf, buf + got);
        fclose(fi);
        jpeg::DecPlan pl;
        if (!jpeg::parse_jpeg(file.data(), file.size(), &pl)) {
            fprintf(stderr, "unsupported: %s\n", pl.error);
            return 6;
        }
        const jpeg::DecFrame &f = pl.f;
        const size_t outb = (size_t)f.H * f.W * (f.ncomp == 3 ? 3 : 1);
        std::vector<unsigned char> out(outb);
        if (std::string(argv[1]) == "--host-decode") {
            std::vector<short> coef(f.nblocks * 64, 0);
            std::vector<unsigned char> planes(f.nsamples);
            const unsigned nmcu = (unsigned)f.mx * f.my, ri = f.restart_interval ? f.restart_interval : nmcu;
            const bool par = argc > 4 && std::string(argv[4]) == "par" && f.nintervals == 1;
            if (par) {  // the self-synchronising decode, serially: same bodies and the same order of passes as the kernels
                const unsigned char *scan = file.data() + pl.scan_begin;
                const size_t nb = f.scan_bytes;
                const unsigned nsub = (unsigned)((nb + jpeg::kSubBytes - 1) / jpeg::kSubBytes);
                std::vector<jpeg::HState> E(nsub), E2(nsub);
                jpeg::NoSink none;
                for (unsigned i = 0; i < nsub; ++i) {
                    jpeg::HState s = {jpeg::sub_start(scan, nb, i), 0, 0, 0, 0};
                    jpeg::decode_span(scan, nb, s, jpeg::sub_limit(scan, nb, i, nsub), f, pl.t, 0, none);
                    E[i] = s;
                }
                int rounds = 0, changed = 1;
                for (; rounds < jpeg::kSyncRounds && changed; ++rounds) {
                    changed = 0;
                    for (unsigned i = 0; i < nsub; ++i) {
                        jpeg::HState s = jpeg::sub_entry(scan, nb, i, E.data());
                        jpeg::decode_span(scan, nb, s, jpeg::sub_limit(scan, nb, i, nsub), f, pl.t, 0, none);
                        if (i + 1 < nsub && !jpeg::same_state(s, E[i])) changed = 1;
                        E2[i] = s;
                    }
                    E.swap(E2);
                }
                fprintf(stderr, "parallel decode: %u subsequences, %d rounds, %s\n", nsub, rounds, changed ? "NOT converged" : "converged");
                if (changed) return 7;
                std::vector<unsigned long long> base(nsub + 1, 0);
                for (unsigned i = 0; i < nsub; ++i) {
                    jpeg::HState s = jpeg::sub_entry(scan, nb, i, E.data());
                    base[i + 1] = base[i] + jpeg::decode_span(scan, nb, s, jpeg::sub_limit(scan, nb, i, nsub), f, pl.t, 0, none);
                }
                const unsigned long long total = (unsigned long long)nmcu * jpeg::blocks_per_mcu(f);
                jpeg::CoefSink sink = {coef.data(), &f, total};
                for (unsigned i = nsub; i-- > 0;) {
                    jpeg::HState s = jpeg::sub_entry(scan, nb, i, E.data());
                    jpeg::decode_span(scan, nb, s, jpeg::sub_limit(scan, nb, i, nsub), f, pl.t, base[i], sink);
                }
                int pred[3] = {0, 0, 0};
                for (unsigned long long o = 0; o < total; ++o) {
                    const int c = jpeg::mcu_component(f, (unsigned)(o % jpeg::blocks_per_mcu(f)));
                    short *b = jpeg::scan_block(coef.data(), f, o);
                    pred[c] += b[0];
                    b[0] = (short)pred[c];
                }
            } else
            for (unsigned k = 0; k < f.nintervals; ++k)
                jpeg::decode_interval(file.data() + pl.scan_begin, (size_t)pl.intervals[2 * k], (size_t)pl.intervals[2 * k + 1], f, pl.t,
                                      k * ri, std::min(k * ri + ri, nmcu), coef.data());
            for (size_t i = 0; i < f.nblocks; ++i) {
                const int c = (f.ncomp == 3 && i >= f.coef_off[2]) ? 2 : (f.ncomp == 3 && i >= f.coef_off[1]) ? 1 : 0;
                const size_t local = i - f.coef_off[c], stride = (size_t)f.bw[c] * 8;
                jpeg::idct_block(coef.data() + i * 64, pl.t.q[f.tq[c]],
                                 planes.data() + f.plane_off[c] + (local / f.bw[c]) * 8 * stride + (local % f.bw[c]) * 8, stride);
            }
            for (int y = 0; y < f.H; ++y)
                for (int x = 0; x < f.W; ++x) jpeg::output_pixel(planes.data(), f, y, x, out.data());
        } else {
            const jpeg::DecWorkspace lay = jpeg::carve_dec(f, nullptr);
            void *d_ws;
            unsigned char *d_out;
            CK(cudaMalloc(&d_ws, lay.bytes));
            CK(cudaMalloc(&d_out, outb));
            CK(cudaMemset(d_ws, 0xA5, lay.bytes));
            const jpeg::DecWorkspace ws = jpeg::carve_dec(f, d_ws);
            cudaEvent_t e0, e1;
            CK(cudaEventCreate(&e0));
            CK(cudaEventCreate(&e1));
            float best = 1e30f;
            for (int rep = 0; rep < 3; ++rep) {
                CK(cudaEventRecord(e0));
                CK(jpeg::decode_launch(file.data(), pl, ws, d_out, 0, nullptr, argc > 4 && std::string(argv[4]) == "par"));
                CK(cudaEventRecord(e1));
                CK(cudaDeviceSynchronize());
                float ms;
                CK(cudaEventElapsedTime(&ms, e0, e1));
                best = ms < best ? ms : best;
            }
            CK(cudaMemcpy(out.data(), d_out, outb, cudaMemcpyDeviceToHost));
            fprintf(stderr, "gpu decode: %dx%d, %u restart intervals, %.3f ms\n", f.H, f.W, f.nintervals, best);
        }
        printf("%d %d %d\n", f.H, f.W, f.ncomp == 3 ? 3 : 1);
        jpeg::free_plan(&pl);
        FILE *fo = fopen(argv[3], "wb");
        if (!fo || fwrite(out.data(), 1, outb, fo) != outb) return 1;
        fclose(fo);
        return 0;
    }
    if (argc < 6) {
        fprintf(stderr, "usage: %s --host|--gpu H W in.rgb out.jpg [quality] [frames]\n", argv[0]);
        return 1;
    }
    const bool gpu = std::string(argv[1]) == "--gpu";
    const int H = atoi(argv[2]), W = atoi(argv[3]), quality = argc > 6 ? atoi(argv[6]) : 75;
    const int n = argc > 7 ? atoi(argv[7]) : 1;   // frames (the same image n times; every copy is checked)
    const size_t px = (size_t)H * W * 3;
    std::vector<uint8_t> rgb(px);
    FILE *f = fopen(argv[4], "rb");
    if (!f || fread(rgb.data(), 1, px, f) != px) {
        fprintf(stderr, "cannot read %zu bytes from %s\n", px, argv[4]);
        return 1;
    }
    fclose(f);
    const size_t cap = px * 2 + 4096;
    std::vector<uint8_t> out(cap);
    size_t nbytes = 0;
    if (!gpu) {
        nbytes = jpeg::encode_host_emulation(rgb.data(), H, W, quality, out.data(), cap);
    } else {
        const jpeg::Geom g = jpeg::make_geom(H, W, 0);
        const jpeg::Workspace lay = jpeg::carve(g, n, nullptr);
        uint8_t *d_rgb, *d_out;
        void *d_ws;
        unsigned long long *d_n;
        CK(cudaMalloc(&d_rgb, px * n));
        CK(cudaMalloc(&d_out, cap * n));
        CK(cudaMalloc(&d_ws, lay.bytes));
        CK(cudaMalloc(&d_n, sizeof(unsigned long long) * n));
        for (int k = 0; k < n; ++k) CK(cudaMemcpy(d_rgb + px * k, rgb.data(), px, cudaMemcpyHostToDevice));
        CK(cudaMemset(d_ws, 0xA5, lay.bytes));  // the encoder must not rely on a clean workspace
        const jpeg::Workspace ws = jpeg::carve(g, n, d_ws);
        int sm = 0;
        CK(cudaDeviceGetAttribute(&sm, cudaDevAttrMultiProcessorCount, 0));
        cudaEvent_t e0, e1;
        CK(cudaEventCreate(&e0));
        CK(cudaEventCreate(&e1));
        float best = 1e30f;
        for (int rep = 0; rep < 3; ++rep) {
            CK(cudaEventRecord(e0));
            CK(jpeg::encode_launch(d_rgb, g, n, quality, ws, d_out, cap, cap, d_n, sm, 0, nullptr));
            CK(cudaEventRecord(e1));
            CK(cudaDeviceSynchronize());
            float ms;
            CK(cudaEventElapsedTime(&ms, e0, e1));
            best = ms < best ? ms : best;
        }
        std::vector<unsigned long long> hn(n);
        CK(cudaMemcpy(hn.data(), d_n, sizeof(unsigned long long) * n, cudaMemcpyDeviceToHost));
        nbytes = (size_t)hn[0];
        if (nbytes > cap) {
            fprintf(stderr, "stream truncated: %zu > %zu\n", nbytes, cap);
            return 3;
        }
        CK(cudaMemcpy(out.data(), d_out, nbytes, cudaMemcpyDeviceToHost));
        std::vector<uint8_t> other(nbytes);
        for (int k = 1; k < n; ++k) {
            if (hn[k] != hn[0]) {
                fprintf(stderr, "frame %d: %llu bytes, frame 0: %llu\n", k, hn[k], hn[0]);
                return 4;
            }
            CK(cudaMemcpy(other.data(), d_out + cap * k, nbytes, cudaMemcpyDeviceToHost));
            if (memcmp(other.data(), out.data(), nbytes)) {
                fprintf(stderr, "frame %d differs from frame 0\n", k);
                return 4;
            }
        }
        printf("gpu: %d x %dx%d -> %zu bytes each, %.3f ms per call (%.1f MP/s)\n", n, H, W, nbytes, best,
               (double)H * W * n / best / 1e3);
    }
    f = fopen(argv[5], "wb");
    if (!f || fwrite(out.data(), 1, nbytes, f) != nbytes) return 1;
    fclose(f);
    return 0;
}
