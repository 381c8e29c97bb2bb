// key_traits.cuh -- the device-side key model.
//
// Replaces RadixDigits::get_digit (reference src/radix_sort/radix_digits.rs).
// The reference re-applies the order transform on every get_digit call
// (radix_digits.rs:59,110-112); here a key is mapped ONCE into its
// order-preserving unsigned image ("ordered" key) by the first executed pass,
// stays in that form between passes, and is mapped back by the last executed
// pass, so the inner loops only extract bytes.
//
//   unsigned  (radix_digits.rs:7-53)    ordered = x
//   signed    (radix_digits.rs:55-101)  ordered = x ^ MIN
//   f32/f64   (radix_digits.rs:103-124) ordered = b ^ ((b >> N-1) | MIN)   (arithmetic >>)
//                                       inverse: t ^ (((t >> N-1) - 1) | MIN) (logical >>)
#pragma once
#include <stdint.h>

namespace b200 {

enum { XF_NONE = 0, XF_SIGNED = 1, XF_FLOAT = 2 };

struct alignas(16) u128 {
    unsigned long long lo, hi;
};

template <int BYTES> struct UIntOf;
template <> struct UIntOf<1> { using type = uint8_t; };
template <> struct UIntOf<2> { using type = uint16_t; };
template <> struct UIntOf<4> { using type = uint32_t; };
template <> struct UIntOf<8> { using type = unsigned long long; };
template <> struct UIntOf<16> { using type = u128; };

// ---- scalar keys (1..8 bytes) -------------------------------------------------
template <typename U, int XF>
__host__ __device__ __forceinline__ U to_ordered(U k) {
    constexpr int BITS = (int)sizeof(U) * 8;
    constexpr U SIGN = (U)((U)1 << (BITS - 1));
    if (XF == XF_SIGNED) return (U)(k ^ SIGN);
    if (XF == XF_FLOAT) {
        U m = (k & SIGN) ? (U)~(U)0 : SIGN;
        return (U)(k ^ m);
    }
    return k;
}
template <typename U, int XF>
__host__ __device__ __forceinline__ U from_ordered(U t) {
    constexpr int BITS = (int)sizeof(U) * 8;
    constexpr U SIGN = (U)((U)1 << (BITS - 1));
    if (XF == XF_SIGNED) return (U)(t ^ SIGN);
    if (XF == XF_FLOAT) {
        U m = (t & SIGN) ? SIGN : (U)~(U)0;
        return (U)(t ^ m);
    }
    return t;
}
template <typename U>
__host__ __device__ __forceinline__ unsigned digit_of(U k, int pass) {
    return (unsigned)(k >> (8 * pass)) & 0xFFu;
}
template <typename U>
__host__ __device__ __forceinline__ bool ordered_less(U a, U b) { return a < b; }
template <typename U>
__host__ __device__ __forceinline__ unsigned long long low64(U k) { return (unsigned long long)k; }

// ---- 128-bit keys (radix_digits.rs:39-45, 87-93) ------------------------------
template <>
__host__ __device__ __forceinline__ u128 to_ordered<u128, XF_NONE>(u128 k) { return k; }
template <>
__host__ __device__ __forceinline__ u128 to_ordered<u128, XF_SIGNED>(u128 k) {
    k.hi ^= 0x8000000000000000ull;
    return k;
}
template <>
__host__ __device__ __forceinline__ u128 from_ordered<u128, XF_NONE>(u128 k) { return k; }
template <>
__host__ __device__ __forceinline__ u128 from_ordered<u128, XF_SIGNED>(u128 k) {
    k.hi ^= 0x8000000000000000ull;
    return k;
}
template <>
__host__ __device__ __forceinline__ unsigned digit_of<u128>(u128 k, int pass) {
    unsigned long long w = pass < 8 ? k.lo : k.hi;
    return (unsigned)(w >> (8 * (pass & 7))) & 0xFFu;
}
template <>
__host__ __device__ __forceinline__ bool ordered_less<u128>(u128 a, u128 b) {
    return a.hi < b.hi || (a.hi == b.hi && a.lo < b.lo);
}
template <>
__host__ __device__ __forceinline__ unsigned long long low64<u128>(u128 k) { return k.lo; }

// prefix_less(a, b, pass): compare two ordered keys on digits 0..pass only (the
// bits above digit `pass` are shifted out).  A tile that a pass has sorted by
// digit `pass` is correctly ordered iff this never holds for (next, current).
template <typename U>
__host__ __device__ __forceinline__ bool prefix_less(U a, U b, int pass) {
    const int sh = (int)sizeof(U) * 8 - 8 * (pass + 1);
    return (U)(a << sh) < (U)(b << sh);
}
template <>
__host__ __device__ __forceinline__ bool prefix_less<u128>(u128 a, u128 b, int pass) {
    if (pass < 8) {
        const int sh = 64 - 8 * (pass + 1);
        return (a.lo << sh) < (b.lo << sh);
    }
    const int sh = 128 - 8 * (pass + 1);
    const unsigned long long ah = a.hi << sh, bh = b.hi << sh;
    return ah < bh || (ah == bh && a.lo < b.lo);
}

// ---- run-time transform selection ---------------------------------------------
// The sort kernels take the transform as a launch parameter (xf) instead of a
// template parameter: it is applied only on the first load and the last store
// of a sort, and one kernel image per key WIDTH keeps the library small.
template <typename U>
__host__ __device__ __forceinline__ U to_ordered_rt(U k, int xf) {
    constexpr int BITS = (int)sizeof(U) * 8;
    constexpr U SIGN = (U)((U)1 << (BITS - 1));
    U m = (xf == XF_NONE) ? (U)0 : SIGN;
    if (xf == XF_FLOAT && (k & SIGN)) m = (U)~(U)0;
    return (U)(k ^ m);
}
template <typename U>
__host__ __device__ __forceinline__ U from_ordered_rt(U t, int xf) {
    constexpr int BITS = (int)sizeof(U) * 8;
    constexpr U SIGN = (U)((U)1 << (BITS - 1));
    U m = (xf == XF_NONE) ? (U)0 : SIGN;
    if (xf == XF_FLOAT && !(t & SIGN)) m = (U)~(U)0;
    return (U)(t ^ m);
}
template <>
__host__ __device__ __forceinline__ u128 to_ordered_rt<u128>(u128 k, int xf) {
    if (xf != XF_NONE) k.hi ^= 0x8000000000000000ull;  // no 128-bit float kind exists
    return k;
}
template <>
__host__ __device__ __forceinline__ u128 from_ordered_rt<u128>(u128 k, int xf) {
    if (xf != XF_NONE) k.hi ^= 0x8000000000000000ull;
    return k;
}

// splitmix64 finaliser: the mixing function of the counter-based generators
// and of the multiset fingerprint.
__host__ __device__ __forceinline__ unsigned long long mix64(unsigned long long z) {
    z += 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

// Key-only element traits: the element moved by the passes IS the key.
template <typename U>
struct KeyTraits {
    using elem_t = U;
    using key_t = U;
    static constexpr int KEY_BYTES = (int)sizeof(U);
    static constexpr int DIGITS = (int)sizeof(U);  // NUMBER_OF_DIGITS, radix_digits.rs:8..117
    // equal keys are bit-identical, so a pass only has to keep same-digit keys in
    // the order of their lower digits -- which K3 can check exactly (sort_kernels.cuh)
    static constexpr bool NEEDS_STABLE = false;
    static constexpr bool FEW_BUCKETS = false;
    struct Params { int xf; };
    __device__ __forceinline__ static bool prefix_less(elem_t a, elem_t b, int pass, const Params&) { return b200::prefix_less<U>(a, b, pass); }
    __device__ __forceinline__ static elem_t xform_in(elem_t e, const Params& p) { return to_ordered_rt<U>(e, p.xf); }
    __device__ __forceinline__ static elem_t xform_out(elem_t e, const Params& p) { return from_ordered_rt<U>(e, p.xf); }
    __device__ __forceinline__ static unsigned digit(elem_t e, int pass, const Params&) { return digit_of<U>(e, pass); }
    // byte address of element 0 of the buffer bucket d is written to
    __device__ __forceinline__ static unsigned long long dst_base(unsigned, const void* dst, const Params&) { return (unsigned long long)dst; }
};

// Partition traits (multi-GPU MSD step): one "pass" whose digit is the
// destination rank of the key, r = #{j : splitter[j] <= ordered(key)}.  Keys are
// moved raw; splitters are held in the ordered domain.
constexpr int MAX_SPLITTERS = 15;
template <typename U>
struct PartitionTraits {
    using elem_t = U;
    using key_t = U;
    static constexpr int KEY_BYTES = (int)sizeof(U);
    static constexpr int DIGITS = 1;
    static constexpr bool NEEDS_STABLE = false;  // order inside a destination is irrelevant: it is sorted next
    // at most 16 buckets and a digit that costs 15 wide compares: K3 computes it once
    // per key (4-bit codes kept in registers) and writes the tile out bucket by bucket
    static constexpr bool FEW_BUCKETS = true;
    struct Params {
        U spl[MAX_SPLITTERS];
        int nsplit;
        int xf;
        // fused exchange: peer[r] != 0 is the receive buffer of rank r as mapped into
        // this process (peer memory over NVLink); bucket r is then stored there
        // directly and bins_in[r] counts elements from the start of THAT buffer
        unsigned long long peer[MAX_SPLITTERS + 1];
        // buckets this launch writes out (bit r = bucket r); the others are ranked but not
        // stored -- the overlapped multi-GPU exchange sends the buckets in two launches
        unsigned mask;
        // claim mode (claim != 0): no counts are known in advance and no look-back runs.  Every tile
        // claims room for its keys of bucket r with ONE system-scope atomicAdd on the cursor of that
        // bucket's region -- cursor[r], a u64 that may live on another GPU -- and stores them at
        // peer[r] + old value (the order inside a region is free: it is sorted next).  A claim that
        // would pass cap[r] elements is not stored and sets bit r of *flags (the caller then takes
        // the exact path); the cursor still counts it, so its final value is the bucket's true size.
        unsigned claim;
        unsigned long long cursor[MAX_SPLITTERS + 1];
        unsigned long long cap[MAX_SPLITTERS + 1];
        unsigned long long flags;
    };
    __device__ __forceinline__ static bool prefix_less(elem_t, elem_t, int, const Params&) { return false; }
    __device__ __forceinline__ static elem_t xform_in(elem_t e, const Params&) { return e; }
    __device__ __forceinline__ static elem_t xform_out(elem_t e, const Params&) { return e; }
    __device__ __forceinline__ static unsigned digit(elem_t e, int, const Params& p) {
        const U k = to_ordered_rt<U>(e, p.xf);
        if (p.nsplit <= 3) {
            // warp-uniform trip count: 2 GPUs cost one compare per key
            unsigned r = 0;
#pragma unroll 1
            for (int j = 0; j < p.nsplit; ++j)
                if (!ordered_less<U>(k, p.spl[j])) ++r;
            return r;
        }
        // 4..15 splitters: branch-free binary search, four compares; the probe of every
        // level is picked by selects among kernel-parameter constants (no indexed
        // loads).  spl[nsplit..14] hold the largest ordered value, so the count can
        // only overshoot for keys equal to it, and is clamped.
        const bool b3 = !ordered_less<U>(k, p.spl[7]);
        U c = b3 ? p.spl[11] : p.spl[3];
        const bool b2 = !ordered_less<U>(k, c);
        const U c0 = b2 ? p.spl[5] : p.spl[1];
        const U c1 = b2 ? p.spl[13] : p.spl[9];
        c = b3 ? c1 : c0;
        const bool b1 = !ordered_less<U>(k, c);
        const U d00 = b1 ? p.spl[2] : p.spl[0];
        const U d01 = b1 ? p.spl[6] : p.spl[4];
        const U d10 = b1 ? p.spl[10] : p.spl[8];
        const U d11 = b1 ? p.spl[14] : p.spl[12];
        const U e0 = b2 ? d01 : d00;
        const U e1 = b2 ? d11 : d10;
        c = b3 ? e1 : e0;
        const bool b0 = !ordered_less<U>(k, c);
        const unsigned r = (b3 ? 8u : 0u) + (b2 ? 4u : 0u) + (b1 ? 2u : 0u) + (b0 ? 1u : 0u);
        return min(r, (unsigned)p.nsplit);
    }
    __device__ __forceinline__ static unsigned long long dst_base(unsigned d, const void* dst, const Params& p) {
        const unsigned long long peer = d <= (unsigned)MAX_SPLITTERS ? p.peer[d] : 0ull;
        return peer ? peer : (unsigned long long)dst;
    }
};

// Record traits: `(K, V)` tuples moved whole, key = one field of the record
// (radix_digits.rs:126-136).  R = unsigned type as wide as the record (8 or 16
// bytes), K = unsigned type of the key, key_shift = 8 * byte offset of the key.
template <typename R, typename K> struct RecordTraits;

template <typename K>
struct RecordTraits<unsigned long long, K> {  // 8-byte record, 4-byte key
    using elem_t = unsigned long long;
    using key_t = K;
    static constexpr int KEY_BYTES = (int)sizeof(K);
    static constexpr int DIGITS = (int)sizeof(K);
    static constexpr bool NEEDS_STABLE = true;  // payloads make the order of equal keys observable (tests.rs:175-187)
    static constexpr bool FEW_BUCKETS = false;
    struct Params { int key_shift; int xf; };
    __device__ __forceinline__ static bool prefix_less(elem_t, elem_t, int, const Params&) { return false; }
    __device__ __forceinline__ static K key(elem_t e, const Params& p) { return (K)(e >> p.key_shift); }
    __device__ __forceinline__ static elem_t with_key(elem_t e, K k, const Params& p) {
        const elem_t mask = (elem_t)(K)~(K)0 << p.key_shift;
        return (e & ~mask) | ((elem_t)k << p.key_shift);
    }
    __device__ __forceinline__ static elem_t xform_in(elem_t e, const Params& p) {
        return with_key(e, to_ordered_rt<K>(key(e, p), p.xf), p);
    }
    __device__ __forceinline__ static elem_t xform_out(elem_t e, const Params& p) {
        return with_key(e, from_ordered_rt<K>(key(e, p), p.xf), p);
    }
    __device__ __forceinline__ static unsigned digit(elem_t e, int pass, const Params& p) {
        return (unsigned)(e >> (p.key_shift + 8 * pass)) & 0xFFu;
    }
    __device__ __forceinline__ static unsigned long long dst_base(unsigned, const void* dst, const Params&) { return (unsigned long long)dst; }
};

template <>
struct RecordTraits<u128, unsigned long long> {  // 16-byte record, 8-byte key
    using elem_t = u128;
    using key_t = unsigned long long;
    static constexpr int KEY_BYTES = 8;
    static constexpr int DIGITS = 8;
    static constexpr bool NEEDS_STABLE = true;
    static constexpr bool FEW_BUCKETS = false;
    struct Params { int key_shift; int xf; };  // key_shift: 0 -> key in .lo, 64 -> key in .hi
    __device__ __forceinline__ static bool prefix_less(elem_t, elem_t, int, const Params&) { return false; }
    __device__ __forceinline__ static key_t key(elem_t e, const Params& p) { return p.key_shift ? e.hi : e.lo; }
    __device__ __forceinline__ static elem_t with_key(elem_t e, key_t k, const Params& p) {
        if (p.key_shift) e.hi = k; else e.lo = k;
        return e;
    }
    __device__ __forceinline__ static elem_t xform_in(elem_t e, const Params& p) {
        return with_key(e, to_ordered_rt<key_t>(key(e, p), p.xf), p);
    }
    __device__ __forceinline__ static elem_t xform_out(elem_t e, const Params& p) {
        return with_key(e, from_ordered_rt<key_t>(key(e, p), p.xf), p);
    }
    __device__ __forceinline__ static unsigned digit(elem_t e, int pass, const Params& p) {
        return (unsigned)(key(e, p) >> (8 * pass)) & 0xFFu;
    }
    __device__ __forceinline__ static unsigned long long dst_base(unsigned, const void* dst, const Params&) { return (unsigned long long)dst; }
};

}  // namespace b200
