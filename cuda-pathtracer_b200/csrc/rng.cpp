// rng.cpp -- host construction of the XORWOW subsequence jump matrices (see rng.h).
//
// T is the 160x160 GF(2) matrix of one xorwow step on (v0..v4) (the Weyl counter d is outside the
// linear part: a jump of n * 2^67 draws adds n * 2^67 * 362437 = 0 mod 2^32 to it).  One cuRAND
// "subsequence" is 2^67 draws, so J[0] = T^(2^67) (67 squarings) and J[b+1] = J[b]^2.
#include <cstring>
#include <vector>

#include "rng.h"

namespace ptx {

namespace {
void square(std::vector<uint32_t>& m, std::vector<uint32_t>& tmp) {
    for (int r = 0; r < kXorwowRows; r++) {
        Xorwow s; const uint32_t* row = &m[(size_t)r * 5];
        s.v0 = row[0]; s.v1 = row[1]; s.v2 = row[2]; s.v3 = row[3]; s.v4 = row[4]; s.d = 0;
        xorwow_apply(s, m.data());
        uint32_t* o = &tmp[(size_t)r * 5];
        o[0] = s.v0; o[1] = s.v1; o[2] = s.v2; o[3] = s.v3; o[4] = s.v4;
    }
    m.swap(tmp);
}
// a := a * b (first a, then b): row r of the product is row r of a pushed through b
void compose(std::vector<uint32_t>& a, const std::vector<uint32_t>& b, std::vector<uint32_t>& tmp) {
    for (int r = 0; r < kXorwowRows; r++) {
        Xorwow s; const uint32_t* row = &a[(size_t)r * 5];
        s.v0 = row[0]; s.v1 = row[1]; s.v2 = row[2]; s.v3 = row[3]; s.v4 = row[4]; s.d = 0;
        xorwow_apply(s, b.data());
        uint32_t* o = &tmp[(size_t)r * 5];
        o[0] = s.v0; o[1] = s.v1; o[2] = s.v2; o[3] = s.v3; o[4] = s.v4;
    }
    a.swap(tmp);
}
void one_step(std::vector<uint32_t>& m) {
    for (int r = 0; r < kXorwowRows; r++) {
        uint32_t e[5] = {0, 0, 0, 0, 0};
        e[r / 32] = 1u << (r % 32);
        Xorwow s; s.v0 = e[0]; s.v1 = e[1]; s.v2 = e[2]; s.v3 = e[3]; s.v4 = e[4]; s.d = 0;
        xorwow_next(s);
        uint32_t* o = &m[(size_t)r * 5];
        o[0] = s.v0; o[1] = s.v1; o[2] = s.v2; o[3] = s.v3; o[4] = s.v4;
    }
}
}  // namespace

// T^n by binary exponentiation (powers of T commute, so the order of the factors does not matter).
void xorwow_build_skip_matrix(uint64_t n, uint32_t* out) {
    std::vector<uint32_t> res((size_t)kXorwowJumpWords, 0u), base((size_t)kXorwowJumpWords), tmp((size_t)kXorwowJumpWords);
    for (int r = 0; r < kXorwowRows; r++) res[(size_t)r * 5 + r / 32] = 1u << (r % 32);      // identity
    one_step(base);
    while (n) {
        if (n & 1ull) compose(res, base, tmp);
        n >>= 1;
        if (n) square(base, tmp);
    }
    memcpy(out, res.data(), sizeof(uint32_t) * kXorwowJumpWords);
}

void xorwow_build_jump_tables(uint32_t* out) {
    std::vector<uint32_t> m((size_t)kXorwowJumpWords), tmp((size_t)kXorwowJumpWords);
    for (int r = 0; r < kXorwowRows; r++) {
        uint32_t e[5] = {0, 0, 0, 0, 0};
        e[r / 32] = 1u << (r % 32);
        Xorwow s; s.v0 = e[0]; s.v1 = e[1]; s.v2 = e[2]; s.v3 = e[3]; s.v4 = e[4]; s.d = 0;
        xorwow_next(s);
        uint32_t* o = &m[(size_t)r * 5];
        o[0] = s.v0; o[1] = s.v1; o[2] = s.v2; o[3] = s.v3; o[4] = s.v4;
    }
    for (int i = 0; i < 67; i++) square(m, tmp);
    for (int b = 0; b < kXorwowJumps; b++) {
        memcpy(out + (size_t)b * kXorwowJumpWords, m.data(), sizeof(uint32_t) * kXorwowJumpWords);
        if (b + 1 < kXorwowJumps) square(m, tmp);
    }
}

}  // namespace ptx
