// host_io.cu — native host-side data plumbing either side of the hot path: no CUDA in this
// file, it is plain C++ compiled into libpcp_b200.so.
//   pcp_crc32c / pcp_tfrecord_scan / pcp_example_decode : the on-disk format of PointCloudSet
//       (pointcloud/PointCloudSet.py:251-288 TFRecordWriter, :306-323 tf.train.Example with the features
//        observations float[n*3], class_label int64, part_labels int64[n], se3 float[9]; schema :100-105)
//   pcp_parse_aftr_frame : the AftrBurner frame text format "(x, y, z) class part" parsed line by line in
//       pointcloud/PointCloudSet.py:161-198 (~47 frames/s in the reference's pure-Python loop)
#include <cmath>
#include <cstdlib>
#include <algorithm>
#include <charconv>
#include <cstring>
#include <vector>

#include "common.cuh"

#if defined(__x86_64__)
#include <nmmintrin.h>
#endif

namespace pcp {

static uint32_t g_crc_table[8][256];
static bool g_crc_init = false;

static void crc_init()
{
    if (g_crc_init) return;
    for (uint32_t i = 0; i < 256; ++i) {
        uint32_t c = i;
        for (int k = 0; k < 8; ++k) c = (c & 1u) ? (c >> 1) ^ 0x82F63B78u : (c >> 1);     // Castagnoli, reflected
        g_crc_table[0][i] = c;
    }
    for (uint32_t i = 0; i < 256; ++i)
        for (int t = 1; t < 8; ++t) g_crc_table[t][i] = (g_crc_table[t - 1][i] >> 8) ^ g_crc_table[0][g_crc_table[t - 1][i] & 0xFFu];
    g_crc_init = true;
}

static uint32_t crc32c_sw(const uint8_t *p, size_t n, uint32_t crc)
{
    crc_init();
    while (n >= 8) {                                   // slicing-by-8
        uint64_t w;
        memcpy(&w, p, 8);
        w ^= crc;
        crc = g_crc_table[7][w & 0xFF] ^ g_crc_table[6][(w >> 8) & 0xFF] ^ g_crc_table[5][(w >> 16) & 0xFF] ^ g_crc_table[4][(w >> 24) & 0xFF] ^
              g_crc_table[3][(w >> 32) & 0xFF] ^ g_crc_table[2][(w >> 40) & 0xFF] ^ g_crc_table[1][(w >> 48) & 0xFF] ^ g_crc_table[0][(w >> 56) & 0xFF];
        p += 8; n -= 8;
    }
    while (n--) crc = (crc >> 8) ^ g_crc_table[0][(crc ^ *p++) & 0xFFu];
    return crc;
}

#if defined(__x86_64__)
__attribute__((target("sse4.2"))) static uint32_t crc32c_hw(const uint8_t *p, size_t n, uint32_t crc)
{
    uint64_t c = crc;
    while (n >= 8) { uint64_t w; memcpy(&w, p, 8); c = _mm_crc32_u64(c, w); p += 8; n -= 8; }
    uint32_t c32 = (uint32_t)c;
    while (n--) c32 = _mm_crc32_u8(c32, *p++);
    return c32;
}
#endif

static uint32_t crc32c(const void *data, size_t n)
{
    const uint8_t *p = reinterpret_cast<const uint8_t *>(data);
#if defined(__x86_64__)
    static const bool hw = __builtin_cpu_supports("sse4.2");
    if (hw) return ~crc32c_hw(p, n, 0xFFFFFFFFu);
#endif
    return ~crc32c_sw(p, n, 0xFFFFFFFFu);
}

static inline uint32_t mask_crc(uint32_t crc) { return ((crc >> 15) | (crc << 17)) + 0xa282ead8u; }    // TFRecord masking

// ---- minimal protobuf wire reader ----
struct Rd {
    const uint8_t *p, *e;
    bool ok = true;
    uint64_t varint()
    {
        uint64_t v = 0;
        for (int s = 0; s < 64 && p < e; s += 7) {
            const uint8_t b = *p++;
            v |= (uint64_t)(b & 0x7F) << s;
            if (!(b & 0x80)) return v;
        }
        ok = false;
        return 0;
    }
    Rd sub()
    {
        const uint64_t n = varint();
        Rd r{p, p, ok};
        if (!ok || n > (uint64_t)(e - p)) { ok = false; r.ok = false; return r; }
        r.e = p + n; p += n;
        return r;
    }
    void skip(uint32_t wt)
    {
        if (wt == 0) varint();
        else if (wt == 1) { if (e - p < 8) ok = false; else p += 8; }
        else if (wt == 2) sub();
        else if (wt == 5) { if (e - p < 4) ok = false; else p += 4; }
        else ok = false;
    }
};

}  // namespace pcp

using namespace pcp;

extern "C" uint32_t pcp_crc32c(const void *h_data, size_t bytes) { return crc32c(h_data, bytes); }
extern "C" uint32_t pcp_crc32c_masked(const void *h_data, size_t bytes) { return mask_crc(crc32c(h_data, bytes)); }

// Walk a TFRecord file image: for each record write the payload offset and length.  verify_crc: 0 none, 1 length CRCs,
// 2 length + payload CRCs.  h_offsets / h_lengths may be NULL (count only).
extern "C" int pcp_tfrecord_scan(const uint8_t *h_buf, size_t bytes, int verify_crc, uint64_t *h_offsets, uint64_t *h_lengths, size_t capacity,
                                 size_t *count)
{
    PCP_REQUIRE(h_buf && count, PCP_ERR_INVALID, "pcp_tfrecord_scan: null argument");
    size_t pos = 0, n = 0;
    while (pos < bytes) {
        PCP_REQUIRE(bytes - pos >= 12, PCP_ERR_INVALID, "pcp_tfrecord_scan: truncated record header at byte %zu", pos);
        uint64_t len; uint32_t lcrc;
        memcpy(&len, h_buf + pos, 8); memcpy(&lcrc, h_buf + pos + 8, 4);
        if (verify_crc >= 1) PCP_REQUIRE(mask_crc(crc32c(h_buf + pos, 8)) == lcrc, PCP_ERR_INVALID, "pcp_tfrecord_scan: corrupt length CRC at byte %zu", pos);
        PCP_REQUIRE(len <= bytes - pos - 12 && bytes - pos - 12 - len >= 4, PCP_ERR_INVALID, "pcp_tfrecord_scan: truncated record payload at byte %zu", pos);
        if (verify_crc >= 2) {
            uint32_t dcrc;
            memcpy(&dcrc, h_buf + pos + 12 + len, 4);
            PCP_REQUIRE(mask_crc(crc32c(h_buf + pos + 12, len)) == dcrc, PCP_ERR_INVALID, "pcp_tfrecord_scan: corrupt payload CRC in record %zu", n);
        }
        if (h_offsets && h_lengths) {
            PCP_REQUIRE(n < capacity, PCP_ERR_INVALID, "pcp_tfrecord_scan: more than %zu records", capacity);
            h_offsets[n] = pos + 12; h_lengths[n] = len;
        }
        ++n;
        pos += 12 + len + 4;
    }
    *count = n;
    return PCP_OK;
}

// Decode one tf.train.Example of the PointCloudSet schema.  Accepts packed and unpacked repeated fields.
extern "C" int pcp_example_decode(const uint8_t *h_rec, size_t bytes, int64_t width, float *h_obs, int64_t *h_class_label, int32_t *h_part_labels,
                                  float *h_se3)
{
    PCP_REQUIRE(h_rec && width >= 1 && h_obs && h_class_label && h_part_labels && h_se3, PCP_ERR_INVALID, "pcp_example_decode: null argument");
    Rd ex{h_rec, h_rec + bytes};
    int got = 0;
    while (ex.ok && ex.p < ex.e) {
        const uint64_t tag = ex.varint();
        if ((tag >> 3) != 1 || (tag & 7) != 2) { ex.skip((uint32_t)(tag & 7)); continue; }
        Rd feats = ex.sub();                                            // Features
        while (feats.ok && feats.p < feats.e) {
            const uint64_t t2 = feats.varint();
            if ((t2 >> 3) != 1 || (t2 & 7) != 2) { feats.skip((uint32_t)(t2 & 7)); continue; }
            Rd entry = feats.sub();                                     // map entry { key = 1, value = 2 }
            const uint8_t *key = nullptr; size_t klen = 0;
            Rd feature{nullptr, nullptr};
            bool have_feature = false;
            while (entry.ok && entry.p < entry.e) {
                const uint64_t t3 = entry.varint();
                if ((t3 & 7) == 2 && (t3 >> 3) == 1) { Rd k = entry.sub(); key = k.p; klen = (size_t)(k.e - k.p); }
                else if ((t3 & 7) == 2 && (t3 >> 3) == 2) { feature = entry.sub(); have_feature = true; }
                else entry.skip((uint32_t)(t3 & 7));
            }
            PCP_REQUIRE(entry.ok, PCP_ERR_INVALID, "pcp_example_decode: malformed feature map entry");
            if (!key || !have_feature) continue;
            auto is = [&](const char *s) { return klen == strlen(s) && memcmp(key, s, klen) == 0; };
            const int which = is("observations") ? 0 : is("class_label") ? 1 : is("part_labels") ? 2 : is("se3") ? 3 : -1;
            if (which < 0) continue;
            const int64_t want = which == 0 ? width * 3 : which == 1 ? 1 : which == 2 ? width : 9;
            int64_t n = 0;
            while (feature.ok && feature.p < feature.e) {
                const uint64_t t4 = feature.varint();
                const uint32_t fld = (uint32_t)(t4 >> 3), wt = (uint32_t)(t4 & 7);
                if (wt != 2) { feature.skip(wt); continue; }
                Rd list = feature.sub();                                // FloatList (2) / Int64List (3)
                const bool is_float = (which == 0 || which == 3);
                PCP_REQUIRE(fld == (is_float ? 2u : 3u), PCP_ERR_INVALID, "pcp_example_decode: feature '%.*s' has the wrong list type", (int)klen, key);
                while (list.ok && list.p < list.e) {
                    const uint64_t t5 = list.varint();
                    if ((t5 >> 3) != 1) { list.skip((uint32_t)(t5 & 7)); continue; }
                    if (is_float) {
                        float *dst = which == 0 ? h_obs : h_se3;
                        if ((t5 & 7) == 2) {                             // packed
                            Rd pk = list.sub();
                            const int64_t cnt = (pk.e - pk.p) / 4;
                            PCP_REQUIRE(pk.ok && n + cnt <= want, PCP_ERR_INVALID, "pcp_example_decode: '%.*s' longer than %lld", (int)klen, key, (long long)want);
                            memcpy(dst + n, pk.p, (size_t)cnt * 4);
                            n += cnt;
                        } else if ((t5 & 7) == 5) {
                            PCP_REQUIRE(list.e - list.p >= 4 && n < want, PCP_ERR_INVALID, "pcp_example_decode: '%.*s' malformed", (int)klen, key);
                            memcpy(dst + n, list.p, 4); list.p += 4; ++n;
                        } else list.skip((uint32_t)(t5 & 7));
                    } else {
                        auto put = [&](uint64_t v) { if (which == 1) *h_class_label = (int64_t)v; else h_part_labels[n] = (int32_t)(int64_t)v; ++n; };
                        if ((t5 & 7) == 2) {
                            Rd pk = list.sub();
                            while (pk.ok && pk.p < pk.e) {
                                const uint64_t v = pk.varint();
                                PCP_REQUIRE(pk.ok && n < want, PCP_ERR_INVALID, "pcp_example_decode: '%.*s' longer than %lld", (int)klen, key, (long long)want);
                                put(v);
                            }
                        } else if ((t5 & 7) == 0) {
                            const uint64_t v = list.varint();
                            PCP_REQUIRE(list.ok && n < want, PCP_ERR_INVALID, "pcp_example_decode: '%.*s' malformed", (int)klen, key);
                            put(v);
                        } else list.skip((uint32_t)(t5 & 7));
                    }
                }
                PCP_REQUIRE(list.ok, PCP_ERR_INVALID, "pcp_example_decode: malformed list in '%.*s'", (int)klen, key);
            }
            PCP_REQUIRE(feature.ok && n == want, PCP_ERR_INVALID, "pcp_example_decode: feature '%.*s' has %lld values, expected %lld", (int)klen, key,
                        (long long)n, (long long)want);
            got |= 1 << which;
        }
        PCP_REQUIRE(feats.ok, PCP_ERR_INVALID, "pcp_example_decode: malformed Features message");
    }
    PCP_REQUIRE(ex.ok, PCP_ERR_INVALID, "pcp_example_decode: malformed Example message");
    PCP_REQUIRE(got == 15, PCP_ERR_INVALID, "pcp_example_decode: missing feature(s), mask %d (need observations, class_label, part_labels, se3)", got);
    return PCP_OK;
}

// Parse one AftrBurner LiDAR frame ("(x, y, z) class part" per line) exactly as PointCloudSet.add_from_aftr_output does
// (pointcloud/PointCloudSet.py:161-198):
//   * position = comma-separated floats between the first '(' and the first ')' of the line;
//   * labels   = the rest of the line split on single spaces, tokens of length <= 1 dropped (:177); there must be exactly
//                two, a known class label then a known part label, otherwise the WHOLE frame is rejected (the reference
//                raises inside the per-frame try block) -> PCP_ERR_INVALID with the reference's message;
//   * a point with a non-finite coordinate is skipped and counted in *h_nonfinite (:188,197-198);
//   * the frame's class label is the label of its last accepted point (:192).
// Labels are passed as '\n'-separated lists (index = position).  h_xyz (capacity x 3 doubles), h_part (capacity).
extern "C" int pcp_parse_aftr_frame(const char *h_text, size_t bytes, const char *class_labels_nl, const char *part_labels_nl, double *h_xyz,
                                    int32_t *h_part, size_t capacity, size_t *count, int32_t *class_label, size_t *h_nonfinite)
{
    PCP_REQUIRE(h_text && class_labels_nl && part_labels_nl && h_xyz && h_part && count && class_label, PCP_ERR_INVALID, "pcp_parse_aftr_frame: null argument");
    auto lookup = [](const char *list, const char *tok, size_t len) -> int {
        int idx = 0;
        const char *s = list;
        while (*s) {
            const char *e = strchr(s, '\n');
            const size_t l = e ? (size_t)(e - s) : strlen(s);
            if (l == len && memcmp(s, tok, len) == 0) return idx;
            ++idx;
            if (!e) break;
            s = e + 1;
        }
        return -1;
    };
    size_t n = 0, bad = 0;
    int32_t cls = -1;
    const char *p = h_text, *end = h_text + bytes;
    while (p < end) {
        const char *nl = reinterpret_cast<const char *>(memchr(p, '\n', (size_t)(end - p)));
        const char *le = nl ? nl : end;
        const char *ls = p;
        p = nl ? nl + 1 : end;
        while (ls < le && (*ls == ' ' || *ls == '\t' || *ls == '\r')) ++ls;            // line.strip()
        while (le > ls && (le[-1] == ' ' || le[-1] == '\t' || le[-1] == '\r')) --le;
        if (ls == le) {                                                               // python: '' -> float('') raises
            set_error("could not convert string to float: ''");
            return PCP_ERR_INVALID;
        }
        const char *lp = reinterpret_cast<const char *>(memchr(ls, '(', (size_t)(le - ls)));
        const char *rp = reinterpret_cast<const char *>(memchr(ls, ')', (size_t)(le - ls)));
        if (!lp || !rp || rp < lp) { set_error("malformed frame line (no '(x, y, z)')"); return PCP_ERR_INVALID; }
        double pos[3];
        int np = 0;
        const char *q = lp + 1;
        while (q <= rp) {
            const char *c = q;
            while (c < rp && *c != ',') ++c;
            double v;
            {
                // fast path: plain decimal / scientific digits (std::from_chars, no locale); everything else float() accepts
                // (sign '+', inf / nan spellings, surrounding blanks) goes through strtod below
                const char *a = q, *z = c;
                while (a < z && *a == ' ') ++a;
                while (z > a && z[-1] == ' ') --z;
                auto fr = std::from_chars(a, z, v);
                if (fr.ec != std::errc() || fr.ptr != z || a == z) {
                    char buf[64];
                    size_t l = (size_t)(c - q);
                    if (l >= sizeof(buf)) { set_error("coordinate token too long"); return PCP_ERR_INVALID; }
                    memcpy(buf, q, l); buf[l] = 0;
                    char *ep = nullptr;
                    v = strtod(buf, &ep);
                    while (ep && (*ep == ' ' || *ep == '\t')) ++ep;
                    if (ep == buf || (ep && *ep)) { set_error("could not convert string to float: '%s'", buf); return PCP_ERR_INVALID; }
                }
            }
            if (np < 3) pos[np] = v;
            ++np;
            q = c + 1;
            if (c == rp) break;
        }
        if (np != 3) { set_error("frame line has %d coordinates, expected 3", np); return PCP_ERR_INVALID; }
        // labels: split on ' ', keep tokens longer than one character
        const char *tok[3]; size_t tlen[3]; int nt = 0;
        q = rp + 1;
        while (q <= le) {
            const char *c = q;
            while (c < le && *c != ' ') ++c;
            if (c - q > 1) {
                if (nt < 3) { tok[nt] = q; tlen[nt] = (size_t)(c - q); }
                ++nt;
            }
            q = c + 1;
        }
        if (nt != 2) {
            char shown[192]; size_t ns = 0; int k = 0;                    // the token list as Python prints it (error path only)
            q = rp + 1;
            while (q <= le) {
                const char *c = q;
                while (c < le && *c != ' ') ++c;
                if (c - q > 1) {
                    const int w = snprintf(shown + ns, sizeof(shown) - ns, "%s'%.*s'", k ? ", " : "", (int)(c - q), q);
                    if (w > 0) ns = ns + (size_t)w < sizeof(shown) ? ns + (size_t)w : sizeof(shown) - 1;
                    ++k;
                }
                q = c + 1;
            }
            shown[ns] = 0;
            set_error("Dataset must contain both a class label and part label. [%s] is not correct.", shown);
            return PCP_ERR_INVALID;
        }
        const int ci = lookup(class_labels_nl, tok[0], tlen[0]);
        if (ci < 0) { set_error("Class label %.*s not known", (int)tlen[0], tok[0]); return PCP_ERR_INVALID; }
        const int pi = lookup(part_labels_nl, tok[1], tlen[1]);
        if (pi < 0) { set_error("Part label %.*s not known", (int)tlen[1], tok[1]); return PCP_ERR_INVALID; }
        if (std::isfinite(pos[0]) && std::isfinite(pos[1]) && std::isfinite(pos[2])) {
            PCP_REQUIRE(n < capacity, PCP_ERR_INVALID, "pcp_parse_aftr_frame: more than %zu points", capacity);
            h_xyz[3 * n] = pos[0]; h_xyz[3 * n + 1] = pos[1]; h_xyz[3 * n + 2] = pos[2];
            h_part[n] = pi;
            cls = ci;
            ++n;
        } else {
            ++bad;
        }
    }
    *count = n;
    *class_label = cls;
    if (h_nonfinite) *h_nonfinite = bad;
    return PCP_OK;
}

// ---- AftrBurner frame writer: Provizio.to_aftr_frame (point_cloud_toolkit/utils/Provizio.py:126-145) ----
// One line per point: "(x, y, z)" followed by " label" for each of the point's labels.  The reference formats every
// coordinate with an f-string, i.e. repr of a Python float — for float32 points too: numpy's scalar __format__ hands the
// value to float.__format__, so a float32 prints with the digits of its DOUBLE value (9.999999747378752e-05 for
// np.float32(1e-4); pinned by tests/golden/aftr_frames.npz, written by the reference's own function).  repr: shortest digits
// that read back to the same double, positional notation for 1e-4 <= |v| < 1e16 (with a trailing ".0" on integers),
// scientific otherwise (mantissa, 'e', sign, at least two exponent digits), "nan" / "inf".
namespace pcp {

template <typename T>
static char *fmt_py_float(char *out, T v)
{
    if (std::isnan(v)) { memcpy(out, "nan", 3); return out + 3; }
    if (std::isinf(v)) { if (v < 0) *out++ = '-'; memcpy(out, "inf", 3); return out + 3; }
    if (v == 0) { if (std::signbit(v)) *out++ = '-'; memcpy(out, "0.0", 3); return out + 3; }
    char sci[64];
    auto r = std::to_chars(sci, sci + sizeof(sci) - 1, v, std::chars_format::scientific);     // shortest round trip: d[.ddd]e[+-]XX
    *r.ptr = 0;
    char *e = sci;
    while (e < r.ptr && *e != 'e') ++e;
    const int exp10 = atoi(e + 1);
    const char *s = sci;
    if (*s == '-') { *out++ = '-'; ++s; }
    char digits[40];
    int nd = 0;
    for (const char *q = s; q < e; ++q) if (*q != '.') digits[nd++] = *q;
    if (exp10 >= -4 && exp10 < 16) {
        if (exp10 < 0) {
            *out++ = '0'; *out++ = '.';
            for (int i = 0; i < -exp10 - 1; ++i) *out++ = '0';
            memcpy(out, digits, nd); out += nd;
        } else {
            for (int i = 0; i <= exp10; ++i) *out++ = i < nd ? digits[i] : '0';
            *out++ = '.';
            if (nd > exp10 + 1) { memcpy(out, digits + exp10 + 1, nd - exp10 - 1); out += nd - exp10 - 1; }
            else *out++ = '0';
        }
    } else {
        *out++ = digits[0];
        if (nd > 1) { *out++ = '.'; memcpy(out, digits + 1, nd - 1); out += nd - 1; }
        *out++ = 'e';
        *out++ = exp10 < 0 ? '-' : '+';
        const int a = exp10 < 0 ? -exp10 : exp10;
        if (a < 10) *out++ = '0';
        auto r2 = std::to_chars(out, out + 8, a);
        out = r2.ptr;
    }
    return out;
}

}  // namespace pcp

// h_points: n x 3, float32 (is_f64 == 0) or float64.  Labels: h_label_ids (n x labels_per_point) int32 indices into the
// '\n'-separated label_table_nl (labels_per_point == 0: none).  Writes the text into h_out (capacity bytes) and its length
// into *len; when the buffer is too small nothing useful is written, *len is the size needed and PCP_ERR_NOMEM is returned.
extern "C" int pcp_format_aftr_frame(const void *h_points, int is_f64, size_t n, const int32_t *h_label_ids, int labels_per_point,
                                     const char *label_table_nl, char *h_out, size_t capacity, size_t *len)
{
    PCP_REQUIRE(len && (n == 0 || h_points) && labels_per_point >= 0 && (labels_per_point == 0 || (h_label_ids && label_table_nl)), PCP_ERR_INVALID,
                "pcp_format_aftr_frame: bad arguments");
    std::vector<std::pair<const char *, size_t>> table;
    size_t longest = 0;
    if (labels_per_point > 0) {
        const char *s = label_table_nl;
        for (;;) {
            const char *e = strchr(s, '\n');
            const size_t l = e ? (size_t)(e - s) : strlen(s);
            table.emplace_back(s, l);
            longest = std::max(longest, l);
            if (!e) break;
            s = e + 1;
        }
    }
    const size_t per_line = 3 * 32 + 8 + (size_t)labels_per_point * (longest + 1);
    const size_t need = n * per_line + 1;
    *len = need;
    if (!h_out || capacity < need) { set_error("pcp_format_aftr_frame: buffer of %zu bytes needed", need); return PCP_ERR_NOMEM; }
    char *o = h_out;
    const float *pf = reinterpret_cast<const float *>(h_points);
    const double *pd = reinterpret_cast<const double *>(h_points);
    for (size_t i = 0; i < n; ++i) {
        *o++ = '(';
        for (int a = 0; a < 3; ++a) {
            o = fmt_py_float(o, is_f64 ? pd[3 * i + a] : (double)pf[3 * i + a]);
            if (a < 2) { *o++ = ','; *o++ = ' '; }
        }
        *o++ = ')';
        for (int k = 0; k < labels_per_point; ++k) {
            const int32_t id = h_label_ids[i * (size_t)labels_per_point + k];
            PCP_REQUIRE(id >= 0 && (size_t)id < table.size(), PCP_ERR_INVALID, "pcp_format_aftr_frame: label id %d out of range at point %zu", id, i);
            *o++ = ' ';
            memcpy(o, table[id].first, table[id].second);
            o += table[id].second;
        }
        *o++ = '\n';
    }
    *len = (size_t)(o - h_out);
    return PCP_OK;
}

// ---- batch decode: the map(parse, num_parallel_calls = AUTOTUNE) + batch() stage of the reference's tf.data pipeline
// (pointcloud/PointCloudSet.py:347-363) — `count` records decoded by `threads` host threads straight into one batch ----
#include <atomic>
#include <thread>

extern "C" int pcp_example_decode_batch(const uint8_t *const *h_recs, const uint64_t *h_lengths, int count, int64_t width, float *h_obs,
                                        int32_t *h_class_labels, int32_t *h_part_labels, float *h_se3, int threads)
{
    PCP_REQUIRE(count >= 0 && width >= 1 && (count == 0 || (h_recs && h_lengths && h_obs && h_class_labels && h_part_labels && h_se3)), PCP_ERR_INVALID,
                "pcp_example_decode_batch: bad arguments");
    if (count == 0) return PCP_OK;
    const int nt = std::max(1, std::min(threads, count));
    std::atomic<int> next{0}, failed{-1};
    std::string first_error;
    std::atomic_flag err_lock = ATOMIC_FLAG_INIT;
    auto work = [&]() {
        for (;;) {
            const int i = next.fetch_add(1);
            if (i >= count || failed.load() >= 0) return;
            int64_t cl = 0;
            const int rc = pcp_example_decode(h_recs[i], (size_t)h_lengths[i], width, h_obs + (size_t)i * width * 3, &cl, h_part_labels + (size_t)i * width,
                                              h_se3 + (size_t)i * 9);
            h_class_labels[i] = (int32_t)cl;
            if (rc != PCP_OK) {
                while (err_lock.test_and_set()) {}
                if (failed.load() < 0) { failed.store(i); first_error = pcp_last_error(); }      // the message is thread-local: carry it over
                err_lock.clear();
            }
        }
    };
    if (nt == 1) work();
    else {
        std::vector<std::thread> pool;
        for (int t = 0; t < nt - 1; ++t) pool.emplace_back(work);
        work();
        for (auto &t : pool) t.join();
    }
    if (failed.load() >= 0) { set_error("record %d of the batch: %s", failed.load(), first_error.c_str()); return PCP_ERR_INVALID; }
    return PCP_OK;
}
