// tests/sanitize/fuzz_host.cpp — TEST HARNESS ONLY.  The host layer's parsers of UNTRUSTED files — the image decoders
// (pupiloptixlab_b200/host/image_io.cpp: PNG / BMP / TGA / PNM / HDR / EXR / PFM written from the format specifications)
// and the scene loader (xml_lite.h + scene_xml.cpp) — built with -fsanitize=address,undefined and driven over seed
// files plus deterministic mutations of them: every truncation length, byte flips, 32-bit words overwritten by extreme
// values (header fields: sizes, offsets, counts).  Any out-of-bounds access, signed overflow or leak-free crash aborts
// the process (-fno-sanitize-recover); the decoders may refuse a file but must never fault.
//
//   fuzz_host <dir>        every regular file in <dir> is a seed; *.xml go to the scene loader, the rest to the decoders
#include <dirent.h>
#include <sys/stat.h>

#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "../../pupiloptixlab_b200/host/image_io.h"
#include "../../pupiloptixlab_b200/host/scene_xml.h"

using namespace pupil::host;

static uint32_t rng_state = 12345u;
static uint32_t rnd() { rng_state = rng_state * 1664525u + 1013904223u; return rng_state >> 8; }

static unsigned long g_runs = 0, g_ok = 0;

static void one(const std::vector<unsigned char>& data, const std::string& name, bool xml, const std::string& dir) {
    ++g_runs;
    std::string err;
    if (xml) {
        Scene sc;
        const std::string text(data.begin(), data.end());
        if (load_scene_xml_text(text, dir, sc, err)) ++g_ok;
    } else {
        Image img;
        // exact-size heap copy: one byte past the end is poisoned for ASan
        std::vector<unsigned char> copy(data);
        copy.shrink_to_fit();
        if (load_image_memory(copy.data(), copy.size(), name, img, err)) {
            ++g_ok;
            if ((size_t)img.w * img.h * 4 != img.rgba.size()) { fprintf(stderr, "%s: size mismatch\n", name.c_str()); abort(); }
        }
    }
}

int main(int argc, char** argv) {
    if (argc < 2) { fprintf(stderr, "usage: fuzz_host <seed dir>\n"); return 2; }
    const std::string dir = argv[1];
    DIR* d = opendir(dir.c_str());
    if (!d) { perror("opendir"); return 2; }
    std::vector<std::string> names;
    while (dirent* e = readdir(d)) {
        const std::string p = dir + "/" + e->d_name;
        struct stat st;
        if (stat(p.c_str(), &st) == 0 && S_ISREG(st.st_mode)) names.push_back(e->d_name);
    }
    closedir(d);
    const uint32_t extremes[] = {0u, 1u, 0x7fffffffu, 0x80000000u, 0xffffffffu, 0x00010000u, 0xffff0000u, 0x01000000u, 65535u, 40000u};
    for (const std::string& nm : names) {
        const std::string p = dir + "/" + nm;
        FILE* f = fopen(p.c_str(), "rb");
        if (!f) continue;
        std::vector<unsigned char> seed;
        unsigned char tmp[4096];
        size_t k;
        while ((k = fread(tmp, 1, sizeof tmp, f)) > 0) seed.insert(seed.end(), tmp, tmp + k);
        fclose(f);
        const bool xml = nm.size() > 4 && nm.substr(nm.size() - 4) == ".xml";
        one(seed, nm, xml, dir);
        // truncations: every length for short files, 256 evenly spread lengths for long ones
        const size_t n = seed.size(), tstep = n <= 512 ? 1 : n / 256;
        for (size_t len = 0; len < n; len += tstep) one(std::vector<unsigned char>(seed.begin(), seed.begin() + len), nm, xml, dir);
        // single byte flips / random bytes
        for (int it = 0; it < 600 && n; ++it) {
            std::vector<unsigned char> m(seed);
            const int cnt = 1 + (int)(rnd() % 3);
            for (int c = 0; c < cnt; ++c) m[rnd() % n] = (unsigned char)rnd();
            one(m, nm, xml, dir);
        }
        // 32-bit header words (first 160 bytes) overwritten by extreme values, both byte orders
        for (size_t off = 0; off + 4 <= n && off < 160; ++off)
            for (uint32_t v : extremes)
                for (int be = 0; be < 2; ++be) {
                    std::vector<unsigned char> m(seed);
                    for (int b = 0; b < 4; ++b) m[off + b] = (unsigned char)(v >> (be ? 24 - 8 * b : 8 * b));
                    one(m, nm, xml, dir);
                }
        // ASCII digit runs replaced by huge numbers (PNM / PFM / HDR headers, XML attributes)
        for (size_t i = 0; i < n && i < 4096; ++i)
            if (seed[i] >= '0' && seed[i] <= '9' && (i == 0 || seed[i - 1] < '0' || seed[i - 1] > '9')) {
                for (const char* big : {"99999999999999999999", "2147483647", "-1", "0", "65536"}) {
                    std::vector<unsigned char> m(seed.begin(), seed.begin() + i);
                    m.insert(m.end(), big, big + strlen(big));
                    size_t j = i;
                    while (j < n && seed[j] >= '0' && seed[j] <= '9') ++j;
                    m.insert(m.end(), seed.begin() + j, seed.end());
                    one(m, nm, xml, dir);
                }
            }
    }
    printf("fuzz_host: %lu inputs from %zu seeds, %lu accepted, no sanitizer report\n", g_runs, names.size(), g_ok);
    return 0;
}
