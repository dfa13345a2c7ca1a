// Row packers: the host side of the packed wire formats of ingestion (psb_ring_store_bitmap / _palette).
#include <stdint.h>
#include <string.h>

#include <vector>

#include <cuda_fp16.h>
#if defined(__SSE2__) && !defined(PSB_NO_SIMD)
#include <emmintrin.h>
#endif

#include "../include/psb200.h"

void psb_set_error(const char* fmt, ...);

// ------------------------------------------------------------------ bitmap packing (host side of psb_ring_store_bitmap)
namespace {
// Bit b of the result: fp16 value p[b] (b < 32) is non-zero; +0.0 and -0.0 are both "absent".
#if defined(__SSE2__) && !defined(PSB_NO_SIMD)
inline uint32_t nonzero_mask32(const uint16_t* p) {
    const __m128i mag = _mm_set1_epi16(0x7fff), zero = _mm_setzero_si128();
    const __m128i a = _mm_cmpeq_epi16(_mm_and_si128(_mm_loadu_si128((const __m128i*)(p)), mag), zero);
    const __m128i b = _mm_cmpeq_epi16(_mm_and_si128(_mm_loadu_si128((const __m128i*)(p + 8)), mag), zero);
    const __m128i c = _mm_cmpeq_epi16(_mm_and_si128(_mm_loadu_si128((const __m128i*)(p + 16)), mag), zero);
    const __m128i d = _mm_cmpeq_epi16(_mm_and_si128(_mm_loadu_si128((const __m128i*)(p + 24)), mag), zero);
    const uint32_t lo = (uint32_t)_mm_movemask_epi8(_mm_packs_epi16(a, b));      // bit set = value is zero
    const uint32_t hi = (uint32_t)_mm_movemask_epi8(_mm_packs_epi16(c, d));
    return ~(lo | (hi << 16));
}
#else
// Four values per 64-bit word: rows are ~95 % zeros, so most words cost one load and one test.
inline uint32_t nonzero_mask32(const uint16_t* p) {
    uint32_t bits = 0;
    for (int q = 0; q < 8; ++q) {
        uint64_t x;
        memcpy(&x, p + 4 * q, 8);
        x &= 0x7fff7fff7fff7fffULL;
        if (x) {
            // lane <= 0x7fff, so lane + 0x7fff sets bit 15 of the lane iff lane != 0 and never carries over;
            // the multiply gathers bits 0 / 16 / 32 / 48 into bits 48..51
            const uint64_t t = ((x + 0x7fff7fff7fff7fffULL) >> 15) & 0x0001000100010001ULL;
            bits |= (uint32_t)(((t * 0x0001000200040008ULL) >> 48) & 0xFu) << (4 * q);
        }
    }
    return bits;
}
#endif

// Appends the values of src[0..32) whose bit is set in `bits` to out[k...] in order and returns the new k.  The first
// four go without a data-dependent branch (a 32-value group of a ~5 % dense row holds 0-3 non-zeros, so a loop over
// the set bits mispredicts its exit on nearly every group); a step with no bit left stores a dummy at out[k], which
// is inside the caller's buffer: such a group has at least 28 absent values, so k <= capacity - 28.
inline long long gather32(const uint16_t* src, uint32_t bits, uint16_t* out, long long k) {
    uint32_t m = bits;
    for (int i = 0; i < 4; ++i) {
        out[k] = src[__builtin_ctz(m | 0x80000000u)];
        k += m != 0;
        m &= m - 1;
    }
    for (; m; m &= m - 1) out[k++] = src[__builtin_ctz(m)];
    return k;
}
}  // namespace

extern "C" long long psb_pack_bitmap_dense(const uint16_t* rows, long long n_rows, int F, uint32_t* out_bitmap,
                                           int* out_row_off, uint16_t* out_values) {
    if (!rows || !out_bitmap || !out_row_off || !out_values || F % 32 != 0) {
        psb_set_error("psb_pack_bitmap_dense: null pointer or num_features not a multiple of 32");
        return -1;
    }
    const int words = F / 32;
    long long k = 0;
    for (long long r = 0; r < n_rows; ++r) {
        const uint16_t* row = rows + (size_t)r * F;
        out_row_off[r] = (int)k;
        for (int w = 0; w < words; ++w) {
            const uint32_t bits = nonzero_mask32(row + w * 32);
            k = gather32(row + w * 32, bits, out_values, k);
            out_bitmap[(size_t)r * words + w] = bits;
        }
    }
    out_row_off[n_rows] = (int)k;
    return k;
}

extern "C" long long psb_pack_bitmap_csr(const long long* indptr, const int* indices, const float* values,
                                         long long n_rows, int F, uint32_t* out_bitmap, int* out_row_off,
                                         uint16_t* out_values) {
    if (!indptr || !indices || !values || !out_bitmap || !out_row_off || !out_values || F % 32 != 0) {
        psb_set_error("psb_pack_bitmap_csr: null pointer or num_features not a multiple of 32");
        return -1;
    }
    const int words = F / 32;
    std::vector<uint16_t> dense((size_t)F, (uint16_t)0);
    std::vector<uint32_t> touched((size_t)words, 0u);
    long long k = 0;
    for (long long r = 0; r < n_rows; ++r) {
        // column order of the source is not trusted (and a column may repeat: the last entry wins): go through a
        // dense scratch row, of which only the 32-value groups this row touched are read and cleared again
        for (long long e = indptr[r]; e < indptr[r + 1]; ++e) {
            const int c = indices[e];
            if (c < 0 || c >= F) continue;
            const __half h = __float2half(values[e]);
            memcpy(&dense[c], &h, 2);
            touched[c >> 5] = 1u;
        }
        out_row_off[r] = (int)k;
        for (int w = 0; w < words; ++w) {
            uint32_t bits = 0;
            if (touched[w]) {
                bits = nonzero_mask32(dense.data() + w * 32);
                k = gather32(dense.data() + w * 32, bits, out_values, k);
                memset(dense.data() + w * 32, 0, 64);
                touched[w] = 0u;
            }
            out_bitmap[(size_t)r * words + w] = bits;
        }
    }
    out_row_off[n_rows] = (int)k;
    return k;
}

// Bitmap + per-row palette: hashed bag-of-words rows hold a handful of distinct values (count / norm, signed), so
// a row ships as its presence bitmap, the distinct fp16 values it uses (<= 16) and one 4-bit code per non-zero —
// bit-exact, a quarter of the value bytes.  A row with more than 16 distinct values ships its fp16 values raw.
// Payload of a row (starts at the even byte offset out_row_off[r]):
//   u16 npal;  npal > 0: u16 palette[npal], then ceil(nnz / 2) code bytes (low nibble first), padded to even
//              npal == 0: u16 values[nnz]
extern "C" long long psb_pack_palette_dense(const uint16_t* rows, long long n_rows, int F, uint32_t* out_bitmap,
                                            int* out_row_off, uint8_t* out_payload, long long payload_capacity) {
    if (!rows || !out_bitmap || !out_row_off || !out_payload || F % 32 != 0) {
        psb_set_error("psb_pack_palette_dense: null pointer or num_features not a multiple of 32");
        return -1;
    }
    const int words = F / 32;
    long long off = 0;
    std::vector<uint8_t> codes((size_t)F + 1);
    std::vector<uint16_t> vals((size_t)F);
    for (long long r = 0; r < n_rows; ++r) {
        const uint16_t* row = rows + (size_t)r * F;
        if (off + 2 + 2LL * F > payload_capacity) {
            psb_set_error("psb_pack_palette_dense: payload buffer too small (row %lld)", r);
            return -1;
        }
        out_row_off[r] = (int)off;
        uint32_t* bitmap = out_bitmap + (size_t)r * words;
        long long nz = 0;                                   // the row's non-zeros, in feature order
        for (int w = 0; w < words; ++w) {
            const uint32_t bits = nonzero_mask32(row + w * 32);
            bitmap[w] = bits;
            nz = gather32(row + w * 32, bits, vals.data(), nz);
        }
        const int nnz = (int)nz;
        alignas(16) uint16_t pal[16] = {0};                 // 0 never equals a non-zero value: free entries do not match
        int npal = 0;
        bool overflow = false;
#if defined(__SSE2__) && !defined(PSB_NO_SIMD)
        __m128i p0 = _mm_setzero_si128(), p1 = _mm_setzero_si128();
        for (int i = 0; i < nnz; ++i) {
            const __m128i h = _mm_set1_epi16((short)vals[i]);
            const int hit = _mm_movemask_epi8(_mm_packs_epi16(_mm_cmpeq_epi16(p0, h), _mm_cmpeq_epi16(p1, h)));
            int c;
            if (hit) {
                c = __builtin_ctz((unsigned)hit);
            } else {
                if (npal == 16) {
                    overflow = true;
                    break;
                }
                pal[npal] = vals[i];
                c = npal++;
                p0 = _mm_load_si128((const __m128i*)pal), p1 = _mm_load_si128((const __m128i*)(pal + 8));
            }
            codes[i] = (uint8_t)c;
        }
#else
        for (int i = 0; i < nnz && !overflow; ++i) {
            const uint16_t h = vals[i];
            int c = 0;
            while (c < npal && pal[c] != h) ++c;
            if (c == npal) {
                if (npal == 16) overflow = true;
                else pal[npal++] = h;
            }
            codes[i] = (uint8_t)c;
        }
#endif
        uint8_t* p = out_payload + off;
        if (overflow || nnz == 0) {
            const uint16_t zero = 0;
            memcpy(p, &zero, 2);
            if (nnz) memcpy(p + 2, vals.data(), (size_t)nnz * 2);
            off += 2 + 2LL * nnz;
        } else {
            const uint16_t np = (uint16_t)npal;
            memcpy(p, &np, 2);
            memcpy(p + 2, pal, (size_t)npal * 2);
            uint8_t* c = p + 2 + 2 * npal;
            const int nb = (nnz + 1) / 2;
            codes[nnz] = 0;                                   // high nibble of the last byte when the count is odd
            for (int i = 0; i < nb; ++i) c[i] = (uint8_t)(codes[2 * i] | (codes[2 * i + 1] << 4));
            off += 2 + 2LL * npal + ((nb + 1) & ~1);
        }
    }
    out_row_off[n_rows] = (int)off;
    return off;
}

