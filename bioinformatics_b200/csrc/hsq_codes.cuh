// hsq_codes.cuh -- device code alphabet and sequence views.
//
// Device code (1 byte): bits 0-3 = nucleotide set (A=1,C=2,G=4,T=8), bit 4 = 'U' marker, bit 5 = GAP.
// Semantics follow COMMON/sequence/IupacNucleotideCode.java:
//   equals  (enum identity, :31-39)  <=> byte equality
//   matches (:185-200: equals, else any shared nucleotide) <=> (a & b & 0x2F) != 0
//     (GAP has an empty nucleotide set and matches only GAP: the private bit 5 gives exactly that;
//      U and T share the T bit; the U marker is masked out of the test.)
#pragma once
#include <stdint.h>

namespace hsq {

static constexpr uint8_t CODE_GAP = 0x20;
static constexpr uint8_t CODE_BAD = 0xFF;
static constexpr uint32_t MATCH_MASK = 0x2F;

// ASCII -> device code (IupacNucleotideCode.getCodesFromString :86-104: upper-cased lookup, '_' '.' '-' = GAP)
__host__ __device__ inline uint8_t encode_ascii(uint8_t ch) {
    switch (ch | 0x20) {  // fold case for letters
        case 'a': return 1;
        case 'c': return 2;
        case 'g': return 4;
        case 't': return 8;
        case 'u': return 0x18;
        case 'r': return 1 | 4;
        case 'y': return 2 | 8;
        case 's': return 4 | 2;
        case 'w': return 1 | 8;
        case 'k': return 4 | 8;
        case 'm': return 1 | 2;
        case 'b': return 2 | 4 | 8;
        case 'd': return 1 | 4 | 8;
        case 'h': return 1 | 2 | 8;
        case 'v': return 1 | 2 | 4;
        case 'n': return 15;
        default: break;
    }
    if (ch == '_' || ch == '.' || ch == '-') return CODE_GAP;
    return CODE_BAD;
}

// device code -> character (toString(): first abbreviation; GAP prints '_')
__host__ __device__ inline uint8_t decode_code(uint8_t c) {
    if (c == 0x18) return 'U';
    if (c & CODE_GAP) return '_';
    const char tbl[17] = "_ACMGRSVTWYHKDBN";
    return (uint8_t)tbl[c & 15];
}

// getComplimentCode (:208-211): swap A<->T, C<->G on the set (= reverse the nibble); U -> A; GAP -> GAP.
__host__ __device__ inline uint8_t complement_code(uint8_t c) {
#ifdef __CUDA_ARCH__
    return (uint8_t)((__brev((uint32_t)c & 15u) >> 28) | (c & CODE_GAP));
#else
    uint32_t m = c & 15u;
    uint32_t r = ((m & 1u) << 3) | ((m & 2u) << 1) | ((m & 4u) >> 1) | ((m & 8u) >> 3);
    return (uint8_t)(r | (c & CODE_GAP));
#endif
}

// A strided, optionally complemented window on a device code array:
// element k = f(base[start + k*step]).  Expresses subSequence / getReverse / getCompliment /
// getReverseCompliment (COMMON/sequence/IupacNucleotideCodeSequence.java:103-117,199-236) without copies.
struct SeqView {
    const uint8_t* base;
    int start;
    int step;
    int len;
    int comp;
    __device__ inline uint8_t get(int k) const {
        uint8_t c = base[start + k * step];
        return comp ? complement_code(c) : c;
    }
};

__device__ inline SeqView make_view(const uint8_t* base, int len, bool reverse, bool comp) {
    SeqView v;
    v.base = base;
    v.len = len;
    v.start = reverse ? len - 1 : 0;
    v.step = reverse ? -1 : 1;
    v.comp = comp ? 1 : 0;
    return v;
}
// sub-window [a, a+n) of a view
__device__ inline SeqView sub_view(const SeqView& s, int a, int n) {
    SeqView v = s;
    v.start = s.start + a * s.step;
    v.len = n;
    return v;
}

}  // namespace hsq
