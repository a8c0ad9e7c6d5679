// pack_core.cuh - the canonical pre-image emitter shared by the host packer (ahv_pack.cpp, multi-threaded C++)
// and the device packer (ahv_pack_dev.cu, one thread per delta): the bytes SemanticDelta.compute_hash hashes
// (audit/delta.py:43-62: json.dumps({...7 keys...}, sort_keys=True).encode()).  One source for both, so the two
// packers cannot drift apart.
//
// Exactness rules reproduced from CPython's json encoder with default arguments:
//   keys sorted at both levels; separators ", " and ": "; None -> null; ints bare;
//   ensure_ascii: '"' -> \", '\' -> \\, \n \r \t \b \f short escapes, other bytes < 0x20 ->
//   \u00XX (lowercase hex), 0x7f -> \u007f, 0x20..0x7e verbatim, non-ASCII code points -> \uXXXX,
//   astral code points -> a UTF-16 surrogate pair of two \uXXXX.
//   timestamp = datetime.isoformat(): YYYY-MM-DDTHH:MM:SS[.ffffff][+HH:MM[:SS]] (the offset part
//   only for aware datetimes).
#pragma once
#include <stdint.h>
#include <string.h>

#if defined(__SSE2__) && !defined(__CUDA_ARCH__)
#include <emmintrin.h>
#define AHV_PACK_SSE2 1
#endif

#include "../../include/ahv.h"

// Everything below is force-inlined into the per-delta loops: the sink is then a local object whose
// address never escapes, so its cursor lives in a register (a byte store through `p` could otherwise alias
// the sink's own members and force a reload per character).
#if defined(__CUDACC__)
#define AHV_PK __host__ __device__ __forceinline__
#else
#define AHV_PK inline __attribute__((always_inline))
#endif

namespace ahv_pack {

template <bool WRITE>
struct SinkT {           // WRITE = false only counts bytes
    uint8_t *p;
    uint64_t n = 0;
    AHV_PK explicit SinkT(uint8_t *dst) : p(dst) {}
    AHV_PK void put(uint8_t c)
    {
        if (WRITE) p[n] = c;
        ++n;
    }
    AHV_PK void lit(const void *s, size_t len)
    {
        if (WRITE) {
#if defined(__CUDA_ARCH__)
            const uint8_t *q = static_cast<const uint8_t *>(s);
            for (size_t k = 0; k < len; ++k) p[n + k] = q[k];
#else
            memcpy(p + n, s, len);
#endif
        }
        n += len;
    }
};
#define LIT(sink, s) (sink).lit(s, sizeof(s) - 1)

AHV_PK uint8_t hexd(uint32_t v) { return (uint8_t)(v < 10 ? '0' + v : 'a' + (v - 10)); }

template <class Sink>
AHV_PK void put_u4(Sink &o, uint32_t cu)   // \uXXXX
{
    o.put('\\'); o.put('u');
    o.put(hexd((cu >> 12) & 15)); o.put(hexd((cu >> 8) & 15)); o.put(hexd((cu >> 4) & 15)); o.put(hexd(cu & 15));
}

// JSON string (with quotes) from UTF-8 bytes, ensure_ascii semantics.  Invalid UTF-8 cannot occur:
// the columns are produced by str.encode().
template <class Sink>
AHV_PK void put_json_string(Sink &o, const uint8_t *s, uint64_t len)
{
    o.put('"');
    for (uint64_t i = 0; i < len;) {
        // run of characters the encoder passes through unchanged
        uint64_t j = i;
#if defined(AHV_PACK_SSE2)
        // 16 bytes at a time: a byte is special when < 0x20, >= 0x7f (also every non-ASCII byte), '"' or '\\'
        while (j + 16 <= len) {
            const __m128i v = _mm_loadu_si128(reinterpret_cast<const __m128i *>(s + j));
            const __m128i lo = _mm_cmplt_epi8(_mm_xor_si128(v, _mm_set1_epi8((char)0x80)), _mm_set1_epi8((char)(0x20 ^ 0x80)));
            const __m128i hi = _mm_cmpgt_epi8(_mm_xor_si128(v, _mm_set1_epi8((char)0x80)), _mm_set1_epi8((char)(0x7e ^ 0x80)));
            const __m128i q = _mm_or_si128(_mm_cmpeq_epi8(v, _mm_set1_epi8('"')), _mm_cmpeq_epi8(v, _mm_set1_epi8('\\')));
            const int m = _mm_movemask_epi8(_mm_or_si128(_mm_or_si128(lo, hi), q));
            if (m) {
                j += (uint64_t)__builtin_ctz((unsigned)m);
                goto run_done;
            }
            j += 16;
        }
#endif
        while (j < len && s[j] >= 0x20 && s[j] < 0x7f && s[j] != '"' && s[j] != '\\') ++j;
#if defined(AHV_PACK_SSE2)
    run_done:
#endif
        if (j > i) {
            o.lit(s + i, j - i);
            i = j;
            continue;
        }
        const uint8_t c = s[i];
        if (c < 0x80) {
            ++i;
            switch (c) {
            case '"': LIT(o, "\\\""); break;
            case '\\': LIT(o, "\\\\"); break;
            case '\n': LIT(o, "\\n"); break;
            case '\r': LIT(o, "\\r"); break;
            case '\t': LIT(o, "\\t"); break;
            case '\b': LIT(o, "\\b"); break;
            case '\f': LIT(o, "\\f"); break;
            default:
                if (c < 0x20 || c == 0x7f) put_u4(o, c);      // the encoder passes only ' '..'~' through
                else o.put(c);
            }
            continue;
        }
        uint32_t cp;
        int extra;
        if ((c & 0xE0) == 0xC0) { cp = c & 0x1F; extra = 1; }
        else if ((c & 0xF0) == 0xE0) { cp = c & 0x0F; extra = 2; }
        else { cp = c & 0x07; extra = 3; }
        ++i;
        for (int k = 0; k < extra && i < len; ++k, ++i) cp = (cp << 6) | (s[i] & 0x3F);
        if (cp >= 0x10000) {
            const uint32_t v = cp - 0x10000;
            put_u4(o, 0xD800 | (v >> 10));
            put_u4(o, 0xDC00 | (v & 0x3FF));
        } else {
            put_u4(o, cp);
        }
    }
    o.put('"');
}

template <class Sink>
AHV_PK void put_int(Sink &o, int64_t v)
{
    char buf[24];
    int k = 0;
    uint64_t u = v < 0 ? (uint64_t)0 - (uint64_t)v : (uint64_t)v;
    do { buf[k++] = (char)('0' + u % 10); u /= 10; } while (u);
    if (v < 0) o.put('-');
    while (k) o.put((uint8_t)buf[--k]);
}

template <class Sink>
AHV_PK void put_2(Sink &o, uint32_t v) { o.put((uint8_t)('0' + v / 10)); o.put((uint8_t)('0' + v % 10)); }

// datetime.isoformat() of (microseconds since the epoch in UTC, utc offset in seconds or INT32_MIN for naive)
template <class Sink>
AHV_PK void put_iso(Sink &o, int64_t epoch_us, int32_t off_s)
{
    const bool aware = off_s != INT32_MIN;
    int64_t local = epoch_us + (aware ? (int64_t)off_s * 1000000 : 0);
    int64_t secs = local / 1000000, us = local % 1000000;
    if (us < 0) { us += 1000000; --secs; }
    int64_t days = secs / 86400, rem = secs % 86400;
    if (rem < 0) { rem += 86400; --days; }
    // civil date from days since 1970-01-01 (proleptic Gregorian)
    int64_t z = days + 719468;
    const int64_t era = (z >= 0 ? z : z - 146096) / 146097;
    const uint32_t doe = (uint32_t)(z - era * 146097);
    const uint32_t yoe = (doe - doe / 1460 + doe / 36524 - doe / 146096) / 365;
    int64_t y = (int64_t)yoe + era * 400;
    const uint32_t doy = doe - (365 * yoe + yoe / 4 - yoe / 100);
    const uint32_t mp = (5 * doy + 2) / 153;
    const uint32_t d = doy - (153 * mp + 2) / 5 + 1;
    const uint32_t m = mp < 10 ? mp + 3 : mp - 9;
    if (m <= 2) ++y;
    o.put((uint8_t)('0' + (y / 1000) % 10)); o.put((uint8_t)('0' + (y / 100) % 10));
    o.put((uint8_t)('0' + (y / 10) % 10)); o.put((uint8_t)('0' + y % 10));
    o.put('-'); put_2(o, m); o.put('-'); put_2(o, d); o.put('T');
    put_2(o, (uint32_t)(rem / 3600)); o.put(':'); put_2(o, (uint32_t)(rem % 3600 / 60)); o.put(':');
    put_2(o, (uint32_t)(rem % 60));
    if (us) {
        o.put('.');
        uint32_t u = (uint32_t)us, div = 100000;
        for (int k = 0; k < 6; ++k) { o.put((uint8_t)('0' + u / div % 10)); div /= 10; }
    }
    if (aware) {
        int32_t a = off_s;
        o.put(a < 0 ? '-' : '+');
        if (a < 0) a = -a;
        put_2(o, (uint32_t)(a / 3600)); o.put(':'); put_2(o, (uint32_t)(a % 3600 / 60));
        if (a % 60) { o.put(':'); put_2(o, (uint32_t)(a % 60)); }
    }
}

struct StrCol {
    const uint8_t *blob;
    const uint64_t *off;     // count+1
    const uint8_t *is_null;  // optional
    template <class Sink>
    AHV_PK void emit(Sink &o, uint64_t i) const
    {
        if (is_null && is_null[i]) LIT(o, "null");
        else put_json_string(o, blob + off[i], off[i + 1] - off[i]);
    }
};

struct Cols {
    uint64_t n;
    StrCol delta_id, session_id, agent_did;
    const int64_t *turn_id, *ts_us;
    const int32_t *ts_off;
    const uint8_t *parent_kind, *parent_digest;
    StrCol parent_other;                 // used where parent_kind == AHV_PARENT_OTHER
    const uint64_t *change_rowptr;
    StrCol ch_path, ch_op, ch_content, ch_prev;
};

// Returns the position, inside the pre-image, of the first character of the parent_hash value
// (past the opening quote when it is a string): what ahv_audit_sessions takes as parent_pos.
template <class Sink>
AHV_PK uint64_t emit_delta(Sink &o, const Cols &c, uint64_t i)
{
    LIT(o, "{\"agent_did\": ");
    c.agent_did.emit(o, i);
    LIT(o, ", \"changes\": [");
    for (uint64_t k = c.change_rowptr[i]; k < c.change_rowptr[i + 1]; ++k) {
        if (k != c.change_rowptr[i]) LIT(o, ", ");
        LIT(o, "{\"content_hash\": ");
        c.ch_content.emit(o, k);
        LIT(o, ", \"operation\": ");
        c.ch_op.emit(o, k);
        LIT(o, ", \"path\": ");
        c.ch_path.emit(o, k);
        LIT(o, ", \"previous_hash\": ");
        c.ch_prev.emit(o, k);
        o.put('}');
    }
    LIT(o, "], \"delta_id\": ");
    c.delta_id.emit(o, i);
    LIT(o, ", \"parent_hash\": ");
    const uint8_t pk = c.parent_kind[i];
    const uint64_t ppos = (pk == AHV_PARENT_NONE) ? 0xffffffffull : o.n + 1;
    if (pk == AHV_PARENT_NONE) LIT(o, "null");
    else if (pk == AHV_PARENT_HEX) {
        o.put('"');
        for (int b = 0; b < 32; ++b) {
            const uint8_t v = c.parent_digest[32 * i + b];
            o.put(hexd(v >> 4)); o.put(hexd(v & 15));
        }
        o.put('"');
    } else c.parent_other.emit(o, i);
    LIT(o, ", \"session_id\": ");
    c.session_id.emit(o, i);
    LIT(o, ", \"timestamp\": \"");
    put_iso(o, c.ts_us[i], c.ts_off[i]);
    LIT(o, "\", \"turn_id\": ");
    put_int(o, c.turn_id[i]);
    o.put('}');
    return ppos;
}

}  // namespace ahv_pack
