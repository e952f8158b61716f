// vr5_volio.cpp - volume files next to the hot path: PVM / DDS / RAW reader and the 16 -> 8 bit
// quantiser, host side (the DDS stream is a serial, data-dependent bit stream). Written from the format's behaviour as the
// reference's loader exhibits it (ddsbase.cpp:187-245 stream, :298-435 containers, :475-558 quantiser; ModelBase.cpp:35-109
// policy), checked byte for byte against that loader on Bucky.pvm and on streams produced by the encoder below.
//
// DDS stream ("DDS v3d\n" / "DDS v3e\n" + payload), bits MSB first:
//   2 bits  skip - 1      byte planes: the decoded stream holds all bytes with index = 0 mod skip first, then 1 mod skip, ...
//   16 bits strip - 1     row length of the 2-D predictor
//   runs: 7 bits count (0 ends the stream), 3 bits width code c (value width = c ? c + 1 : 0), then `count` values of that
//         width, each a residual biased by half the range. Predictor: previous byte, plus (byte above - byte above-left)
//         once more than `strip` bytes exist; everything modulo 256.
//   v3e un-planes in blocks of skip * 2^24 bytes, v3d over the whole stream.
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <string>
#include <vector>

#include "../../include/vr5.h"

namespace {

const char kMagicD[] = "DDS v3d\n", kMagicE[] = "DDS v3e\n";
const uint32_t kPlaneBlock = 1u << 24;

struct BitReader {
    const uint8_t *p;
    size_t n, pos = 0;  // pos in bits
    BitReader(const uint8_t *data, size_t bytes) : p(data), n(bytes) {}
    uint32_t get(unsigned bits) {  // bits <= 16; reads past the end yield zeros, as a zero-padded stream would
        uint32_t v = 0;
        for (unsigned i = 0; i < bits; i++, pos++) {
            const size_t byte = pos >> 3;
            const uint32_t bit = byte < n ? (p[byte] >> (7 - (pos & 7))) & 1u : 0u;
            v = (v << 1) | bit;
        }
        return v;
    }
};

struct BitWriter {
    std::vector<uint8_t> out;
    unsigned fill = 0;
    void put(uint32_t v, unsigned bits) {
        for (int i = (int)bits - 1; i >= 0; i--) {
            if (fill == 0) out.push_back(0);
            out.back() |= (uint8_t)(((v >> i) & 1u) << (7 - fill));
            fill = (fill + 1) & 7;
        }
    }
};

// planes -> interleaved bytes (decoder side), per block of skip * block bytes (block == 0: one block)
void unplane(std::vector<uint8_t> &d, uint32_t skip, uint32_t block) {
    if (skip <= 1 || d.empty()) return;
    const size_t chunk = block ? (size_t)skip * block : d.size();
    std::vector<uint8_t> tmp;
    for (size_t base = 0; base < d.size(); base += chunk) {
        const size_t len = d.size() - base < chunk ? d.size() - base : chunk;
        tmp.assign(d.begin() + base, d.begin() + base + len);
        size_t src = 0;
        for (uint32_t plane = 0; plane < skip; plane++)
            for (size_t j = plane; j < len; j += skip) d[base + j] = tmp[src++];
    }
}

// interleaved bytes -> planes (encoder side)
void plane(std::vector<uint8_t> &d, uint32_t skip, uint32_t block) {
    if (skip <= 1 || d.empty()) return;
    const size_t chunk = block ? (size_t)skip * block : d.size();
    std::vector<uint8_t> tmp;
    for (size_t base = 0; base < d.size(); base += chunk) {
        const size_t len = d.size() - base < chunk ? d.size() - base : chunk;
        tmp.assign(d.begin() + base, d.begin() + base + len);
        size_t dst = 0;
        for (uint32_t pl = 0; pl < skip; pl++)
            for (size_t j = pl; j < len; j += skip) d[base + dst++] = tmp[j];
    }
}

inline int predictor(const std::vector<uint8_t> &out, size_t cnt, uint32_t strip) {
    return (strip == 1 || cnt <= strip) ? 0 : (int)out[cnt - strip] - (int)out[cnt - strip - 1];
}

bool dds_decode(const uint8_t *payload, size_t bytes, uint32_t block, std::vector<uint8_t> &out) {
    BitReader br(payload, bytes);
    const uint32_t skip = br.get(2) + 1, strip = br.get(16) + 1;
    out.clear();
    int act = 0;
    for (;;) {
        const uint32_t run = br.get(7);
        if (run == 0) break;
        const uint32_t code = br.get(3);
        const unsigned width = code ? code + 1 : 0;
        const int bias = (1 << width) / 2;
        for (uint32_t i = 0; i < run; i++) {
            act += predictor(out, out.size(), strip) + (int)br.get(width) - bias;
            act &= 255;  // modulo 256 (the value never strays further than one wrap)
            out.push_back((uint8_t)act);
        }
        if (br.pos > (bytes + 8) * 8) return false;  // ran off the end of a truncated stream
    }
    unplane(out, skip, block);
    return true;
}

// Encoder (tests, tools): runs of up to 127 residuals, each run as wide as its largest residual needs.
void dds_encode(const uint8_t *data, size_t bytes, uint32_t skip, uint32_t strip, uint32_t block, std::vector<uint8_t> &stream) {
    std::vector<uint8_t> planes(data, data + bytes);
    plane(planes, skip, block);
    BitWriter bw;
    bw.put(skip - 1, 2);
    bw.put(strip - 1, 16);
    std::vector<uint8_t> done;
    done.reserve(bytes);
    std::vector<int> resid(bytes);
    int act = 0;
    for (size_t i = 0; i < bytes; i++) {  // residual in [-128, 127] of each byte against its prediction
        done.push_back(planes[i]);
        const int pred = (act + predictor(done, i, strip)) & 255;
        int r = ((int)planes[i] - pred) & 255;
        if (r >= 128) r -= 256;
        resid[i] = r;
        act = planes[i];
    }
    for (size_t i = 0; i < bytes;) {
        const size_t run = bytes - i < 127 ? bytes - i : 127;
        unsigned width = 0;
        for (size_t j = i; j < i + run; j++)
            for (;;) {  // smallest legal width (0, 2..8) whose biased range holds the residual
                const int bias = (1 << width) / 2;
                if (resid[j] >= -bias && resid[j] < (1 << width) - bias) break;
                width = width == 0 ? 2 : width + 1;
            }
        bw.put((uint32_t)run, 7);
        bw.put(width ? width - 1 : 0, 3);
        const int bias = (1 << width) / 2;
        for (size_t j = i; j < i + run; j++) bw.put((uint32_t)(resid[j] + bias), width);
        i += run;
    }
    bw.put(0, 7);
    stream.assign(kMagicD, kMagicD + 8);
    if (block) stream.assign(kMagicE, kMagicE + 8);
    stream.insert(stream.end(), bw.out.begin(), bw.out.end());
}

bool read_file(const char *path, std::vector<uint8_t> &buf) {
    FILE *f = fopen(path, "rb");
    if (!f) return false;
    uint8_t tmp[1 << 16];
    size_t got;
    buf.clear();
    while ((got = fread(tmp, 1, sizeof(tmp), f)) > 0) buf.insert(buf.end(), tmp, tmp + got);
    fclose(f);
    return true;
}

thread_local std::string g_io_error;
int io_fail(const std::string &msg) { g_io_error = msg; return 1; }

}  // namespace

// ---- containers -----------------------------------------------------------------------------------------------------
extern "C" const char *vr5_volume_io_error(void) { return g_io_error.c_str(); }

extern "C" void vr5_volume_free(void *p) { free(p); }

// readDDSfile + readRAWfile fallback (ddsbase.cpp:298-340, :360-361): a file that does not start with a DDS magic is taken as is
static bool load_maybe_dds(const char *path, std::vector<uint8_t> &payload) {
    std::vector<uint8_t> file;
    if (!read_file(path, file)) return false;
    if (file.size() >= 8 && (!memcmp(file.data(), kMagicD, 8) || !memcmp(file.data(), kMagicE, 8))) {
        const uint32_t block = memcmp(file.data(), kMagicE, 8) ? 0u : kPlaneBlock;
        return dds_decode(file.data() + 8, file.size() - 8, block, payload);
    }
    payload.swap(file);
    return true;
}

extern "C" int vr5_dds_decode(const uint8_t *stream, uint64_t bytes, uint8_t **out, uint64_t *out_bytes) {
    if (!stream || !out || !out_bytes || bytes < 8) return io_fail("vr5_dds_decode: bad arguments");
    const bool v3d = !memcmp(stream, kMagicD, 8), v3e = !memcmp(stream, kMagicE, 8);
    if (!v3d && !v3e) return io_fail("vr5_dds_decode: no DDS magic");
    std::vector<uint8_t> payload;
    if (!dds_decode(stream + 8, bytes - 8, v3e ? kPlaneBlock : 0u, payload)) return io_fail("vr5_dds_decode: truncated stream");
    *out = (uint8_t *)malloc(payload.size() ? payload.size() : 1);
    if (!*out) return io_fail("out of memory");
    memcpy(*out, payload.data(), payload.size());
    *out_bytes = payload.size();
    return 0;
}

extern "C" int vr5_dds_encode(const uint8_t *data, uint64_t bytes, uint32_t skip, uint32_t strip, int v3e, uint8_t **out, uint64_t *out_bytes) {
    if (!data || !out || !out_bytes || skip < 1 || skip > 4 || strip < 1 || strip > 65536) return io_fail("vr5_dds_encode: bad arguments");
    std::vector<uint8_t> stream;
    dds_encode(data, bytes, skip, strip, v3e ? kPlaneBlock : 0u, stream);
    *out = (uint8_t *)malloc(stream.size());
    if (!*out) return io_fail("out of memory");
    memcpy(*out, stream.data(), stream.size());
    *out_bytes = stream.size();
    return 0;
}

// next text line of the header; returns false at the end of the buffer
static bool next_line(const std::vector<uint8_t> &b, size_t &pos, std::string &line) {
    if (pos >= b.size()) return false;
    size_t e = pos;
    while (e < b.size() && b[e] != '\n') e++;
    line.assign((const char *)b.data() + pos, e - pos);
    pos = e < b.size() ? e + 1 : e;
    return true;
}

// readPVMvolume (ddsbase.cpp:343-435): "PVM\n" (comment lines start with #), "PVM2\n" / "PVM3\n" (+ voxel scales; PVM3
// appends four NUL-terminated strings after the voxels). Returns the voxels only.
extern "C" int vr5_load_pvm(const char *path, uint8_t **data, uint32_t *width, uint32_t *height, uint32_t *depth, uint32_t *components,
                            float *scale_xyz) {
    if (!path || !data || !width || !height || !depth || !components) return io_fail("vr5_load_pvm: bad arguments");
    std::vector<uint8_t> b;
    if (!load_maybe_dds(path, b)) return io_fail(std::string("vr5_load_pvm: cannot read or decode ") + path);
    size_t pos = 0;
    std::string line;
    if (!next_line(b, pos, line)) return io_fail("vr5_load_pvm: empty file");
    int version = line == "PVM" ? 1 : line == "PVM2" ? 2 : line == "PVM3" ? 3 : 0;
    if (!version) return io_fail("vr5_load_pvm: not a PVM volume");
    unsigned w = 0, h = 0, d = 0, nc = 0;
    float sx = 1.0f, sy = 1.0f, sz = 1.0f;
    do {
        if (!next_line(b, pos, line)) return io_fail("vr5_load_pvm: truncated header");
    } while (version == 1 && !line.empty() && line[0] == '#');
    if (sscanf(line.c_str(), "%u %u %u", &w, &h, &d) != 3 || w < 1 || h < 1 || d < 1) return io_fail("vr5_load_pvm: bad dimensions");
    if (version > 1) {
        if (!next_line(b, pos, line) || sscanf(line.c_str(), "%g %g %g", &sx, &sy, &sz) != 3 || sx <= 0 || sy <= 0 || sz <= 0) return io_fail("vr5_load_pvm: bad voxel scale");
    }
    if (!next_line(b, pos, line) || sscanf(line.c_str(), "%u", &nc) != 1 || nc < 1) return io_fail("vr5_load_pvm: bad component count");
    // the four header fields are untrusted: bound them before multiplying (a product that wraps modulo 2^64 would pass the size
    // check below with a tiny buffer); dimensions beyond 65535 cannot be rendered anyway (ushort3 dims, ModelBase.h:14)
    if (w > 65535u || h > 65535u || d > 65535u || nc > 4u) return io_fail("vr5_load_pvm: dimensions above 65535 or more than 4 components");
    const uint64_t voxels = (uint64_t)w * h * d * nc;   // < 2^50
    if (b.size() - pos < voxels) return io_fail("vr5_load_pvm: voxel data shorter than the header says");
    if (version < 3 && b.size() - pos != voxels) return io_fail("vr5_load_pvm: trailing bytes after the voxel data");
    *data = (uint8_t *)malloc(voxels);
    if (!*data) return io_fail("out of memory");
    memcpy(*data, b.data() + pos, voxels);
    *width = w; *height = h; *depth = d; *components = nc;
    if (scale_xyz) { scale_xyz[0] = sx; scale_xyz[1] = sy; scale_xyz[2] = sz; }
    return 0;
}

// ModelBase::load_model's RAW branch (ModelBase.cpp:66-84) without the interactive prompt: dimensions come from the caller
extern "C" int vr5_load_raw(const char *path, uint32_t width, uint32_t height, uint32_t depth, uint32_t components, uint8_t **data) {
    if (!path || !data || !width || !height || !depth || !components) return io_fail("vr5_load_raw: bad arguments");
    std::vector<uint8_t> b;
    if (!read_file(path, b)) return io_fail(std::string("vr5_load_raw: cannot read ") + path);
    if (b.size() != (uint64_t)width * height * depth * components) return io_fail("vr5_load_raw: file size differs from width*height*depth*components");
    *data = (uint8_t *)malloc(b.size());
    if (!*data) return io_fail("out of memory");
    memcpy(*data, b.data(), b.size());
    return 0;
}

// quantize() (ddsbase.cpp:475-558): 16-bit big-endian voxels -> 8 bit. linear: 255 * v / vmax. Otherwise a non-linear map
// that spends the 256 output levels where the data has structure: weight[v] = cbrt(sum of sqrt(gradient magnitude) over the
// voxels of value v), weights of the extreme values zeroed, peaks clipped to 1/256 of the total (up to 256 passes), then the
// cumulative weight rescaled to 0..255. All in double precision, same operation order, round half up at the end.
extern "C" int vr5_quantize16(const uint8_t *data16_be, uint32_t width, uint32_t height, uint32_t depth, int linear, uint8_t **data8) {
    if (!data16_be || !data8 || !width || !height || !depth) return io_fail("vr5_quantize16: bad arguments");
    const size_t n = (size_t)width * height * depth;
    std::vector<uint16_t> v(n);
    int vmin = 65535, vmax = 0;
    for (size_t i = 0; i < n; i++) {
        const int x = 256 * data16_be[2 * i] + data16_be[2 * i + 1];
        v[i] = (uint16_t)x;
        if (x < vmin) vmin = x;
        if (x > vmax) vmax = x;
    }
    std::vector<double> map(65536, 0.0);
    if (linear) {
        for (int i = 0; i < 65536; i++) map[i] = 255 * (double)i / vmax;
    } else {
        auto at = [&](uint32_t i, uint32_t j, uint32_t k) { return (int)v[i + ((size_t)j + (size_t)k * height) * width]; };
        auto axis = [&](int lo, int mid, int hi, bool has_lo, bool has_hi) -> double {  // central difference, one-sided at the faces
            if (has_lo) return has_hi ? (hi - lo) / 2.0 : (double)(mid - lo);
            return has_hi ? (double)(hi - mid) : 0.0;
        };
        for (uint32_t k = 0; k < depth; k++)
            for (uint32_t j = 0; j < height; j++)
                for (uint32_t i = 0; i < width; i++) {
                    const int c = at(i, j, k);
                    const double gx = axis(i > 0 ? at(i - 1, j, k) : 0, c, i < width - 1 ? at(i + 1, j, k) : 0, i > 0, i < width - 1);
                    const double gy = axis(j > 0 ? at(i, j - 1, k) : 0, c, j < height - 1 ? at(i, j + 1, k) : 0, j > 0, j < height - 1);
                    const double gz = axis(k > 0 ? at(i, j, k - 1) : 0, c, k < depth - 1 ? at(i, j, k + 1) : 0, k > 0, k < depth - 1);
                    map[c] += sqrt(sqrt(gx * gx + gy * gy + gz * gz));
                }
        for (int i = 0; i < 65536; i++) map[i] = pow(map[i], 1.0 / 3);
        map[vmin] = map[vmax] = 0.0;
        for (int pass = 0; pass < 256; pass++) {
            double total = 0.0;
            for (int i = 0; i < 65536; i++) total += map[i];
            bool done = true;
            for (int i = 0; i < 65536; i++)
                if (map[i] > total / 256) { map[i] = total / 256; done = false; }
            if (done) break;
        }
        for (int i = 1; i < 65536; i++) map[i] += map[i - 1];
        if (map[65535] > 0.0f)
            for (int i = 0; i < 65536; i++) map[i] *= 255.0f / map[65535];
    }
    *data8 = (uint8_t *)malloc(n);
    if (!*data8) return io_fail("out of memory");
    for (size_t i = 0; i < n; i++) (*data8)[i] = (uint8_t)(int)(map[v[i]] + 0.5);
    return 0;
}

// ModelBase::load_model (ModelBase.cpp:35-109) for .pvm: decode, quantise two-component data, refuse more than two
extern "C" int vr5_load_volume(const char *path, uint8_t **data8, uint32_t *width, uint32_t *height, uint32_t *depth) {
    uint8_t *raw = nullptr;
    uint32_t nc = 0;
    if (vr5_load_pvm(path, &raw, width, height, depth, &nc, nullptr)) return 1;
    if (nc > 2) { free(raw); return io_fail("vr5_load_volume: unsupported number of components (1 or 2 allowed)"); }
    if (nc == 2) {
        uint8_t *q = nullptr;
        const int rc = vr5_quantize16(raw, *width, *height, *depth, 0, &q);
        free(raw);
        if (rc) return 1;
        raw = q;
    }
    *data8 = raw;
    return 0;
}
