// CSV tokenizer shared by the bulk loader (host_runtime.cpp) and the producer thread (host_producer.cpp).
// Reference: CsvProducer.java:46-58 (split on ',', drop zeros, last column = label).
#pragma once
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include <string>

namespace psb_host {
// Eight ASCII digits (first character in the lowest byte; bytes already reduced to 0..9) -> their integer.
inline uint64_t digits8(uint64_t x) {
    x = (x * 2561) >> 8;
    x = ((x & 0x00FF00FF00FF00FFULL) * 6553601) >> 16;
    return ((x & 0x0000FFFF0000FFFFULL) * 42949672960001ULL) >> 32;
}
// High bit of every byte of `t` (= text ^ "00000000") that is NOT a digit.
inline uint64_t non_digits(uint64_t t) {
    return (((t & 0x7f7f7f7f7f7f7f7fULL) + 0x7676767676767676ULL) | t) & 0x8080808080808080ULL;
}
// The common field "[-]d.ffffff," (one integer digit, up to 15 fraction digits): the fraction digits are found and
// converted eight at a time instead of one character per loop trip.  On success *v is the value, computed exactly
// as the general path computes it (integer mantissa / exact power of ten), and the ',' is returned; nullptr sends
// the field to the general path.
inline const char* fast_decimal(const char* s, const char* e, const double* pow10, float* v) {
    const bool neg = *s == '-';
    s += neg;
    if (e - s < 18) return nullptr;                    // two 8-byte loads behind "d."
    const unsigned d0 = (unsigned)(unsigned char)s[0] - (unsigned)'0';
    if (d0 > 9u || s[1] != '.') return nullptr;
    uint64_t a, b;
    memcpy(&a, s + 2, 8);
    a ^= 0x3030303030303030ULL;
    uint64_t nd = non_digits(a), mant;
    int n;
    if (nd) {
        n = __builtin_ctzll(nd) >> 3;                  // 0..7 fraction digits
        if (s[2 + n] != ',') return nullptr;
        mant = d0;
        if (n) mant = mant * (uint64_t)pow10[n] + digits8(a << (8 * (8 - n)));
    } else {
        memcpy(&b, s + 10, 8);
        b ^= 0x3030303030303030ULL;
        nd = non_digits(b);
        if (!nd) return nullptr;                       // 16 or more fraction digits
        const int n2 = __builtin_ctzll(nd) >> 3;       // 0..7 more
        if (s[10 + n2] != ',') return nullptr;
        mant = d0 * 100000000ULL + digits8(a);
        if (n2) mant = mant * (uint64_t)pow10[n2] + digits8(b << (8 * (8 - n2)));
        n = 8 + n2;
    }
    const float r = (float)((double)mant / pow10[n]);
    *v = neg ? -r : r;
    return s + 2 + n;
}

// Parses one line [p, e): calls emit(col, value) for every non-zero feature, returns the label.
// One pass per field: "0," and "0.0," (nearly every field of a hashed bag-of-words row) are skipped outright,
// plain decimals accumulate an integer mantissa that is divided by an exact power of ten, anything else
// (exponents, more than 18 digits, inf / nan) goes through strtof.
template <class Emit>
inline int parse_line(const char* p, const char* e, int* ncols, Emit emit) {
    static const double kPow10[23] = {1e0,  1e1,  1e2,  1e3,  1e4,  1e5,  1e6,  1e7,  1e8,  1e9,  1e10, 1e11,
                                      1e12, 1e13, 1e14, 1e15, 1e16, 1e17, 1e18, 1e19, 1e20, 1e21, 1e22};
    int col = 0;
    float last = 0.f;
    while (p < e) {
        if (*p == '0') {
            while (e - p >= 8) {               // runs of "0,0,0,0,": four absent features per 64-bit compare
                uint64_t x;
                memcpy(&x, p, 8);
                x ^= 0x2c302c302c302c30ULL;
                if (x) {                       // jump over the "0," pairs in front of the first byte that differs
                    const int k = (__builtin_ctzll(x) >> 3) & ~1;
                    p += k, col += k >> 1;
                    break;
                }
                p += 8, col += 4;
            }
            if (p >= e) break;
            if (*p != '0') continue;
            if (p + 1 < e && p[1] == ',') {
                p += 2, ++col;
                continue;
            }
            if (p + 3 < e && p[1] == '.' && p[2] == '0' && p[3] == ',') {
                p += 4, ++col;
                continue;
            }
        }
        float v;
        const char* s = fast_decimal(p, e, kPow10, &v);
        if (s) {                   // (always a feature column: a ',' follows)
            if (v != 0.f) emit(col, v);
            ++col;
            p = s + 1;
            continue;
        }
        s = p;
        bool neg = false, simple = true, frac = false;
        if (*s == '-' || *s == '+') neg = *s == '-', ++s;
        uint64_t mant = 0;
        int digits = 0, frac_digits = 0;
        for (; s < e; ++s) {
            const unsigned d = (unsigned)(unsigned char)*s - (unsigned)'0';
            if (d <= 9u) {
                mant = mant * 10u + d;
                ++digits, frac_digits += frac;
            } else if (*s == ',') {
                break;
            } else if (*s == '.' && !frac) {
                frac = true;
            } else if (*s == ' ' || *s == '\r') {
            } else {
                simple = false;
                break;
            }
        }
        if (simple && digits <= 18 && frac_digits <= 22) {
            v = (float)((double)mant / kPow10[frac_digits]);
            if (neg) v = -v;
        } else {
            const char* q = (const char*)memchr(p, ',', (size_t)(e - p));
            if (!q) q = e;
            v = strtof(std::string(p, q).c_str(), nullptr);
            s = q;
        }
        if (s < e) {               // a feature column
            if (v != 0.f) emit(col, v);
        } else {
            last = v;              // last column: the label
        }
        ++col;
        p = s + 1;
    }
    *ncols = col;
    return last >= -2147483648.f && last < 2147483648.f ? (int)last : 0;      // (nan / inf / huge labels: 0, not UB)
}
}  // namespace psb_host
