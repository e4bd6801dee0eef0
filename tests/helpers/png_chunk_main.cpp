#include "rct_png.h"
#include <cstdio>
int main(int argc, char **argv) {
    FILE *f = fopen(argv[1], "rb"); std::vector<uint8_t> d; uint8_t b[4096]; size_t k;
    while ((k = fread(b, 1, sizeof b, f)) > 0) d.insert(d.end(), b, b + k);
    fclose(f);
    rct_png::Info info;
    if (!rct_png::parse_header(d.data(), d.size(), &info)) { printf("ERR %s\n", info.error.c_str()); return 1; }
    std::vector<uint8_t> rgb((size_t)info.W * info.H * 3);
    if (!rct_png::decode(d.data(), d.size(), &info, rgb.data())) { printf("ERR %s\n", info.error.c_str()); return 1; }
    fwrite(rgb.data(), 1, rgb.size(), stdout); return 0;
}
