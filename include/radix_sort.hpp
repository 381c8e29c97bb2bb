// radix_sort.hpp -- C++ mirror of the reference's sort interface over the C ABI
// (include/b200sort.h).  Header-only; C++17.
//
// The reference is a Rust crate whose whole public surface for this path is a
// trait method on slices:
//
//     pub trait RadixDigits { const NUMBER_OF_DIGITS: u8; fn get_digit(&self, u8) -> u8; }
//                                                        (radix_digits.rs:1-5)
//     pub trait RadixSort<T: RadixDigits> { fn radix_sort(&mut self); }   (mod.rs:18-20)
//     impl<T: RadixDigits> RadixSort<T> for [T]                          (mod.rs:61)
//
// No Rust toolchain exists in this environment (INTEGRATION.md shows the shim a
// maintainer would add), so this is the compiled-language host side that can be
// built and run here: the same names, the same argument meaning, the same error
// behaviour, one level above the C ABI.
//
//     std::vector<std::uint64_t> v = ...;
//     radix_sort::radix_sort(v);                      // data_sorted.radix_sort()
//     std::vector<std::pair<uint32_t, uint32_t>> r;   // Vec<(u32, u32)>
//     radix_sort::radix_sort(r);                      // stable by .first
//
// RadixDigits<T> is the trait: NUMBER_OF_DIGITS as in the reference, KIND naming
// the built-in impl (radix_digits.rs:7-124).  get_digit itself runs on the GPU;
// a user-defined RadixDigits impl cannot, so an unspecialised T does not compile
// (the C ABI's B200SORT_ERR_UNSUPPORTED_KIND, decided at compile time here).
// Failure is a panic in the reference (`expect`, mod.rs:68,106); here it is an
// exception of type radix_sort::Error carrying b200sort_last_error().
#ifndef RADIX_SORT_HPP
#define RADIX_SORT_HPP

#include <cstddef>
#include <cstdint>
#include <stdexcept>
#include <string>
#include <type_traits>
#include <utility>
#include <vector>

#include "b200sort.h"

namespace radix_sort {

struct Error : std::runtime_error {
    int code;
    Error(int c, const std::string& what) : std::runtime_error("b200sort error " + std::to_string(c) + ": " + what), code(c) {}
};

inline void check(int rc) {
    if (rc != B200SORT_OK) throw Error(rc, b200sort_last_error());
}

// ---- RadixDigits (radix_digits.rs:1-124) ---------------------------------------
template <class T, class = void>
struct RadixDigits;  // no definition: T has no built-in get_digit

#define RADIX_SORT_DIGITS(T, K)                                   \
    template <>                                                   \
    struct RadixDigits<T, void> {                                 \
        static constexpr unsigned NUMBER_OF_DIGITS = sizeof(T);   \
        static constexpr int KIND = K;                            \
    };
RADIX_SORT_DIGITS(std::uint8_t, B200SORT_U8)    // radix_digits.rs:7-13
RADIX_SORT_DIGITS(std::uint16_t, B200SORT_U16)  // :15-21
RADIX_SORT_DIGITS(std::uint32_t, B200SORT_U32)  // :23-29
RADIX_SORT_DIGITS(std::uint64_t, B200SORT_U64)  // :31-37 (and usize, :47-53)
RADIX_SORT_DIGITS(std::int8_t, B200SORT_I8)     // :55-61
RADIX_SORT_DIGITS(std::int16_t, B200SORT_I16)   // :63-69
RADIX_SORT_DIGITS(std::int32_t, B200SORT_I32)   // :71-77
RADIX_SORT_DIGITS(std::int64_t, B200SORT_I64)   // :79-85 (and isize, :95-101)
RADIX_SORT_DIGITS(float, B200SORT_F32)          // :106-114
RADIX_SORT_DIGITS(double, B200SORT_F64)         // :116-124
#if defined(__SIZEOF_INT128__)
RADIX_SORT_DIGITS(unsigned __int128, B200SORT_U128)  // :39-45
RADIX_SORT_DIGITS(__int128, B200SORT_I128)           // :87-93
#endif
#undef RADIX_SORT_DIGITS

// `long long` / `unsigned long long` where they are types of their own (LP64: int64_t is long)
template <class T>
struct RadixDigits<T, std::enable_if_t<(std::is_same_v<T, long long> && !std::is_same_v<long long, std::int64_t>) ||
                                       (std::is_same_v<T, unsigned long long> && !std::is_same_v<unsigned long long, std::uint64_t>)>> {
    static constexpr unsigned NUMBER_OF_DIGITS = sizeof(T);
    static constexpr int KIND = std::is_signed_v<T> ? B200SORT_I64 : B200SORT_U64;
};

// (T, U): the key is field 0, the whole element moves (radix_digits.rs:126-136)
template <class K, class V>
struct RadixDigits<std::pair<K, V>, void> {
    static constexpr unsigned NUMBER_OF_DIGITS = RadixDigits<K>::NUMBER_OF_DIGITS;
    static constexpr int KIND = RadixDigits<K>::KIND;
};

// ---- RadixSort (mod.rs:18-20,61): in place, blocking ------------------------------
template <class T>
inline void radix_sort(T* data, std::size_t len) {
    check(b200sort_keys_host(data, len, RadixDigits<T>::KIND));
}

// records: stable by the key field, like the reference's LSD passes over whole elements
template <class K, class V>
inline void radix_sort(std::pair<K, V>* data, std::size_t len) {
    using R = std::pair<K, V>;
    static_assert(std::is_trivially_copyable_v<K> && std::is_trivially_copyable_v<V>, "elements are moved bit for bit");
    // the two record shapes the library sorts (b200sort_api.cu, check_record_shape); anything else fails to compile
    // here instead of returning B200SORT_ERR_UNSUPPORTED_KIND at run time where the reference would have sorted
    static_assert((sizeof(R) == 8 && sizeof(K) == 4) || (sizeof(R) == 16 && sizeof(K) == 8),
                  "libb200sort sorts (K, V) records of 8 bytes with a 4-byte key or of 16 bytes with an 8-byte key only");
    // layout, not assumed: a Rust tuple is repr(Rust) and a std::pair has padding rules of its own
    const R probe{};
    const std::size_t key_off =
        static_cast<std::size_t>(reinterpret_cast<const unsigned char*>(&probe.first) - reinterpret_cast<const unsigned char*>(&probe));
    check(b200sort_records_host(data, len, sizeof(R), key_off, RadixDigits<K>::KIND));
}

template <class T, class A>
inline void radix_sort(std::vector<T, A>& data) {
    radix_sort(data.data(), data.size());
}

// RadixSortCopyOnly (mod.rs:22-38) and counting_sort (mod.rs:40-59): the reference's
// experimental variants of the same sort; here they are the same call
template <class T, class A> inline void radix_sort0(std::vector<T, A>& d) { radix_sort(d); }
template <class T, class A> inline void radix_sort1(std::vector<T, A>& d) { radix_sort(d); }
template <class T, class A> inline void radix_sort2(std::vector<T, A>& d) { radix_sort(d); }
template <class T, class A> inline void radix_sort3(std::vector<T, A>& d) { radix_sort(d); }
template <class T, class A> inline void radix_sort4(std::vector<T, A>& d) { radix_sort(d); }
template <class T, class A> inline void radix_sort5(std::vector<T, A>& d) { radix_sort(d); }
inline void counting_sort(std::vector<std::uint8_t>& d) { radix_sort(d); }

}  // namespace radix_sort
#endif  // RADIX_SORT_HPP
