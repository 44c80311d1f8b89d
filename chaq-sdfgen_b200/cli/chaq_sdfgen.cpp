// chaq_sdfgen -- host program with the command line of the reference's OpenMP tool (reference
// openmp/sdfgen.c:32-49 usage, :139-244 argument handling, :250-258 load, :305-347 store), running the signed
// distance field on a B200 through the C ABI of include/chaqsdf.h (chaqsdf_u8 replaces sdfgen.c:261-300).
//
// Kept from the reference, because scripts depend on it:
//   * options are recognised by the character after '-' only; "-i -" / "-o -" mean stdin / stdout;
//   * -i -o -s -q -f take the next argument; any other "-xyz" is a bundle of flags out of h, n, l, a, and
//     unknown letters are ignored;  -h prints the usage and exits 0;
//   * spread 0, quality outside 1..100, or a missing -i / -o: usage on stdout, message on stderr, exit status -1;
//   * output type: -f wins, else the three characters after the last '.' of the output name compared to
//     png bmp jpg tga, else PNG;  the result is a 1-channel 8-bit image;
//   * the image is decoded to (luminance, alpha) pairs; the alpha channel is tested unless -l is given.
// Added: -t prints the time spent in the library call to stderr (the reference ignores unknown flags); -g n cuts the
// image into row bands over the first n GPUs of the box (chaqsdf_u8_multi; same bytes); binary PGM (P5, maxval 255)
// is read and written directly (input recognised by its magic, output by "-f pgm" or a .pgm name), for images
// beyond what stb_image can decode -- a PGM has no alpha, so like any image without alpha it is all "inside" unless
// -l tests its grey values. A PGM with a lower maxval goes to stb_image like every other format (as in the reference);
// a 16-bit PGM is refused: stb_image 2.27 -- the reference's decoder -- overruns its buffer converting it
// (stb_image.h:1180 via :1743), so the reference itself crashes on such a file.
//
// Image decoding/encoding is stb_image / stb_image_write, compiled from the reference's include directory
// (they are its third-party submodule; nothing is copied into this repository).
#include <chrono>
#include <climits>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#define STB_IMAGE_IMPLEMENTATION
#define STB_IMAGE_WRITE_IMPLEMENTATION
#include "stb/stb_image.h"
#include "stb/stb_image_write.h"

#include "chaqsdf.h"

namespace {

enum class Kind { Unknown, Png, Bmp, Jpg, Tga, Pgm };

struct Options {
    const char* in = nullptr;
    const char* out = nullptr;
    size_t spread = 64, quality = 100, gpus = 1;
    size_t channel = 1;  // 1 = alpha, 0 = luminance
    bool above = true, asymmetric = false, timing = false;
    bool from_stdin = false, to_stdout = false;
    Kind kind = Kind::Unknown;
};

void print_usage() {
    std::puts("usage: chaq_sdfgen [-f filetype] -i file -o file [-q n] [-s n] [-ahln] [-g n] [-t]\n"
              "    -f filetype   output type: png, bmp, jpg, tga (or pgm) (default: from the output name, else png)\n"
              "    -i file       input image; \"-\" reads stdin\n"
              "    -o file       output image; \"-\" writes stdout\n"
              "    -q n          jpeg quality, 1-100 (default 100)\n"
              "    -s n          spread radius in pixels (default 64)\n"
              "    -a            asymmetric spread: negative distances are dropped (default symmetric)\n"
              "    -h            this text\n"
              "    -l            test the luminance instead of the alpha channel\n"
              "    -n            inverted test: values below the threshold are inside\n"
              "    -g n          row bands over the first n GPUs of this box (default 1)\n"
              "    -t            print the time of the distance-field call to stderr");
}

[[noreturn]] void die(bool with_usage, const char* fmt, ...) {
    if (with_usage) print_usage();
    va_list ap;
    va_start(ap, fmt);
    std::vfprintf(stderr, fmt, ap);
    va_end(ap);
    std::fputc('\n', stderr);
    std::exit(-1);
}

Kind kind_of(const char* s) {
    static const struct { const char* tag; Kind k; } table[] = {{"png", Kind::Png}, {"bmp", Kind::Bmp}, {"jpg", Kind::Jpg}, {"tga", Kind::Tga}, {"pgm", Kind::Pgm}};
    for (const auto& e : table)
        if (std::strncmp(s, e.tag, 3) == 0) return e.k;
    return Kind::Unknown;
}

// returns false when -h asked for the usage (exit 0)
bool parse(int argc, char** argv, Options* o) {
    for (int i = 0; i < argc; ++i) {
        const char* a = argv[i];
        if (a[0] != '-') continue;
        const char key = a[1];
        const bool takes_value = key == 'i' || key == 'o' || key == 's' || key == 'q' || key == 'f' || key == 'g';
        if (takes_value) {
            const char* val = (i + 1 < argc) ? argv[++i] : nullptr;
            switch (key) {
                case 'i':
                    if (!val) { if (!o->in) die(true, "No input file specified."); break; }
                    o->from_stdin = std::strcmp(val, "-") == 0;
                    o->in = val;
                    break;
                case 'o':
                    if (!val) { if (!o->out) die(true, "No output file specified."); break; }
                    o->to_stdout = std::strcmp(val, "-") == 0;
                    o->out = val;
                    break;
                case 's':
                    if (!val) die(true, "No number specified with spread.");
                    o->spread = std::strtoull(val, nullptr, 10);
                    break;
                case 'q':
                    if (!val) die(true, "No number specified with quality.");
                    o->quality = std::strtoull(val, nullptr, 10);
                    break;
                case 'g':
                    if (!val) die(true, "No number specified with gpus.");
                    o->gpus = std::strtoull(val, nullptr, 10);
                    if (o->gpus == 0 || o->gpus > 64) die(true, "Invalid value given for gpus. Must be between 1-64");
                    break;
                default:  // 'f'
                    if (!val) die(true, "Filetype not specified with filetype switch.");
                    o->kind = kind_of(val);
                    if (o->kind == Kind::Unknown) die(true, "Invalid filetype specified.");
                    break;
            }
            continue;
        }
        for (const char* f = a + 1; *f; ++f) {
            if (*f == 'h') { print_usage(); return false; }
            if (*f == 'n') o->above = false;
            else if (*f == 'l') o->channel = 0;
            else if (*f == 'a') o->asymmetric = true;
            else if (*f == 't') o->timing = true;
        }
    }
    if (o->quality == 0 || o->quality > 100) die(true, "Invalid value given for jpeg quality. Must be between 1-100");
    if (o->spread == 0) die(true, "Invalid value given for spread. Must be a positive integer.");
    if (!o->in) die(true, "No input file specified.");
    if (!o->out) die(true, "No output file specified.");
    return true;
}

void to_stdout(void*, void* data, int size) { std::fwrite(data, (size_t)size, 1, stdout); }

// a byte source: a file, or a stream already in memory (stdin cannot be rewound)
struct Src {
    FILE* f = nullptr;
    const unsigned char* p = nullptr;
    size_t n = 0, i = 0;
    int get() { return f ? std::fgetc(f) : (i < n ? p[i++] : EOF); }
    size_t read(unsigned char* dst, size_t cnt) {
        if (f) return std::fread(dst, 1, cnt, f);
        const size_t k = cnt < n - i ? cnt : n - i;
        std::memcpy(dst, p + i, k);
        i += k;
        return k;
    }
};

// binary PGM: "P5" <w> <h> <maxval> <one whitespace> bytes; comments (#...) allowed between header tokens.
// The source stands right after the magic. 0: not a header; 1: 8-bit (maxval 255); 2: a lower maxval (stb's job);
// 3: 16-bit samples.
int read_pgm_header(Src& in, size_t* w, size_t* h) {
    size_t v[3];
    for (int i = 0; i < 3; ++i) {
        int c = in.get();
        while (c == ' ' || c == '\t' || c == '\n' || c == '\r' || c == '#') {
            if (c == '#') while (c != '\n' && c != EOF) c = in.get();
            c = in.get();
        }
        if (c < '0' || c > '9') return 0;
        size_t x = 0;
        while (c >= '0' && c <= '9') { x = x * 10 + (size_t)(c - '0'); if (x > ((size_t)1 << 40)) return 0; c = in.get(); }
        v[i] = x;
        if (i == 2 && !(c == ' ' || c == '\t' || c == '\n' || c == '\r')) return 0;
    }
    *w = v[0]; *h = v[1];
    if (v[0] == 0 || v[1] == 0) return 0;
    return v[2] == 255 ? 1 : (v[2] < 255 ? 2 : 3);
}

}  // namespace

int main(int argc, char** argv) {
    Options o;
    if (!parse(argc, argv, &o)) return 0;

    int w = 0, h = 0, n = 0;
    unsigned char* ya = nullptr;
    std::vector<unsigned char> grey;  // PGM input: one byte per pixel
    std::vector<unsigned char> all;   // stdin: the whole stream
    Src in;
    if (o.from_stdin) {
        unsigned char buf[1 << 16];
        size_t got;
        while ((got = std::fread(buf, 1, sizeof buf, stdin)) > 0) all.insert(all.end(), buf, buf + got);
        in.p = all.data(); in.n = all.size();
    } else {
        in.f = std::fopen(o.in, "rb");
        if (!in.f) die(false, "Input file could not be opened.");
    }
    bool is_pgm8 = false;
    if (in.get() == 'P' && in.get() == '5') {
        size_t pw = 0, ph = 0;
        const int kind = read_pgm_header(in, &pw, &ph);
        if (kind == 3) die(false, "16-bit PGM input is not supported (stb_image 2.27 overruns its buffer on it).");
        if (kind == 1) {
            if (pw > 32768 || ph > 32768)
                die(false, "Input image is %zu x %zu: sides above 32768 are beyond this build's limit (CHAQSDF_MAX_DIM).", pw, ph);
            grey.resize(pw * ph);
            if (in.read(grey.data(), grey.size()) != grey.size()) die(false, "Input file could not be opened.");
            if (o.channel == 1) std::memset(grey.data(), 255, grey.size());  // no alpha channel: opaque everywhere
            w = (int)pw; h = (int)ph;
            is_pgm8 = true;
        }
    }
    if (in.f) std::fclose(in.f);
    if (!is_pgm8) {  // every other format, including PGMs that are not 8-bit: stb_image, as the reference does
        if (o.from_stdin) {
            if (all.size() > (size_t)INT_MAX) die(false, "Input file could not be opened.");
            ya = stbi_load_from_memory(all.data(), (int)all.size(), &w, &h, &n, 2);
        } else {
            ya = stbi_load(o.in, &w, &h, &n, 2);
        }
        if (!ya) die(false, "Input file could not be opened.");
        if (w > 32768 || h > 32768)
            die(false, "Input image is %d x %d: sides above 32768 are beyond this build's limit (CHAQSDF_MAX_DIM).", w, h);
    }
    std::vector<unsigned char>().swap(all);

    std::vector<unsigned char> sdf((size_t)w * (size_t)h);
    const auto t0 = std::chrono::steady_clock::now();
    const unsigned char* src = ya ? ya : grey.data();
    const size_t stride = ya ? 2 : 1, chan = ya ? o.channel : 0;
    int rc;
    if (o.gpus > 1) {
        std::vector<int> devs(o.gpus);
        for (size_t k = 0; k < o.gpus; ++k) devs[k] = (int)k;
        rc = chaqsdf_u8_multi(src, (size_t)w, (size_t)h, stride, chan, o.above ? 1 : 0, o.spread, o.asymmetric ? 1 : 0, devs.data(),
                              (int)devs.size(), sdf.data());
    } else {
        rc = chaqsdf_u8(src, (size_t)w, (size_t)h, stride, chan, o.above ? 1 : 0, o.spread, o.asymmetric ? 1 : 0, sdf.data());
    }
    const double ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    if (ya) stbi_image_free(ya);
    if (rc != CHAQSDF_OK) die(false, "distance field failed (%d): %s", rc, chaqsdf_last_error_string());
    if (o.timing)
        std::fprintf(stderr, "chaqsdf_u8: %dx%d spread %zu on %zu GPU(s) in %.3f ms (%.1f Mpix/s, host buffers)\n", w, h, o.spread, o.gpus, ms,
                     w * (double)h / ms / 1e3);

    if (!o.to_stdout && o.kind == Kind::Unknown) {
        if (const char* dot = std::strrchr(o.out, '.')) o.kind = kind_of(dot + 1);
    }
    int ok = 0;
    switch (o.kind) {
        case Kind::Bmp:
            ok = o.to_stdout ? stbi_write_bmp_to_func(to_stdout, nullptr, w, h, 1, sdf.data()) : stbi_write_bmp(o.out, w, h, 1, sdf.data());
            break;
        case Kind::Jpg:
            ok = o.to_stdout ? stbi_write_jpg_to_func(to_stdout, nullptr, w, h, 1, sdf.data(), (int)o.quality)
                             : stbi_write_jpg(o.out, w, h, 1, sdf.data(), (int)o.quality);
            break;
        case Kind::Pgm: {
            FILE* fo = o.to_stdout ? stdout : std::fopen(o.out, "wb");
            if (fo) {
                std::fprintf(fo, "P5\n%d %d\n255\n", w, h);
                ok = std::fwrite(sdf.data(), 1, sdf.size(), fo) == sdf.size();
                if (fo != stdout) std::fclose(fo);
            }
            break;
        }
        case Kind::Tga:
            ok = o.to_stdout ? stbi_write_tga_to_func(to_stdout, nullptr, w, h, 1, sdf.data()) : stbi_write_tga(o.out, w, h, 1, sdf.data());
            break;
        default:
            ok = o.to_stdout ? stbi_write_png_to_func(to_stdout, nullptr, w, h, 1, sdf.data(), w) : stbi_write_png(o.out, w, h, 1, sdf.data(), w);
            break;
    }
    (void)ok;  // like the reference, a failed write is not an error status
    return 0;
}
