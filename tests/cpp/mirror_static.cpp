// mirror_static.cpp -- compile-time facts of include/radix_sort.hpp against the
// reference's RadixDigits impls (radix_digits.rs:7-136).  Compiling it IS the test.
#include <string>

#include "radix_sort.hpp"

using radix_sort::RadixDigits;

template <class T, class = void> struct has_digits : std::false_type {};
template <class T> struct has_digits<T, std::void_t<decltype(RadixDigits<T>::KIND)>> : std::true_type {};

#define DIGITS(T, N, K)                                                                        \
    static_assert(RadixDigits<T>::NUMBER_OF_DIGITS == N, #T ": NUMBER_OF_DIGITS");               \
    static_assert(RadixDigits<T>::KIND == K, #T ": kind");                                     \
    static_assert(b200sort_kind(RadixDigits<T>::KIND) < B200SORT_KIND_COUNT, #T ": kind id")

DIGITS(std::uint8_t, 1, B200SORT_U8);            // radix_digits.rs:8
DIGITS(std::uint16_t, 2, B200SORT_U16);          // :16
DIGITS(std::uint32_t, 4, B200SORT_U32);          // :24
DIGITS(std::uint64_t, 8, B200SORT_U64);          // :32
DIGITS(unsigned __int128, 16, B200SORT_U128);    // :40
DIGITS(std::size_t, 8, B200SORT_U64);            // :48 (usize, 64-bit target)
DIGITS(std::int8_t, 1, B200SORT_I8);             // :56
DIGITS(std::int16_t, 2, B200SORT_I16);           // :64
DIGITS(std::int32_t, 4, B200SORT_I32);           // :72
DIGITS(std::int64_t, 8, B200SORT_I64);           // :80
DIGITS(__int128, 16, B200SORT_I128);             // :88
DIGITS(std::ptrdiff_t, 8, B200SORT_I64);         // :96 (isize)
DIGITS(long long, 8, B200SORT_I64);
DIGITS(unsigned long long, 8, B200SORT_U64);
DIGITS(float, 4, B200SORT_F32);                  // :107
DIGITS(double, 8, B200SORT_F64);                 // :117

// (T, U): digits and kind of field 0 (radix_digits.rs:126-136)
using P32 = std::pair<std::uint32_t, std::uint32_t>;
using P64 = std::pair<std::uint64_t, std::uint64_t>;
static_assert(RadixDigits<P32>::NUMBER_OF_DIGITS == 4 && RadixDigits<P32>::KIND == B200SORT_U32, "(u32, u32)");
static_assert(RadixDigits<P64>::NUMBER_OF_DIGITS == 8 && RadixDigits<P64>::KIND == B200SORT_U64, "(u64, u64)");
static_assert(sizeof(P32) == 8 && sizeof(P64) == 16, "record shapes the C ABI takes");

// a type without a built-in get_digit has no RadixDigits: sorting it does not compile
static_assert(has_digits<std::uint32_t>::value, "detector works");
static_assert(!has_digits<std::string>::value, "std::string has no get_digit");
static_assert(!has_digits<bool>::value && !has_digits<char>::value, "bool / char are not key types of the reference");
static_assert(!has_digits<long double>::value, "no f128 / long double impl");

int main() { return 0; }
