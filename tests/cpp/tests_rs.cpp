// tests_rs.cpp -- the reference's test suite (src/radix_sort/tests.rs, 15 tests)
// restated one for one in C++ over include/radix_sort.hpp, the compiled-language
// mirror of `slice.radix_sort()`.  Same sizes (1e6), same value families, same
// comparators (slice::sort, sort_by(total_cmp), stable sort_by_key); the only
// liberty is a fixed seed per test where the reference draws from thread_rng.
// Like libtest, the tests run on parallel threads.  Exit status = failed tests.
#include <algorithm>
#include <atomic>
#include <cstdio>
#include <cstring>
#include <functional>
#include <limits>
#include <mutex>
#include <random>
#include <string>
#include <thread>
#include <vector>

#include "radix_sort.hpp"



namespace {

constexpr std::size_t N = 1000000;  // `1e6 as usize`, tests.rs:27,36,...

struct NotSorted : std::runtime_error {
    NotSorted() : std::runtime_error("Not sorted properly!") {}  // tests.rs:14,20
};

// rand::thread_rng().fill(slice) (tests.rs:28): uniform bits
template <class T>
std::vector<T> random_bits(std::uint64_t seed) {
    std::mt19937_64 rng(seed);
    std::vector<T> v(N);
    unsigned char* p = reinterpret_cast<unsigned char*>(v.data());
    const std::size_t bytes = N * sizeof(T);
    for (std::size_t i = 0; i < bytes; i += 8) {
        const std::uint64_t x = rng();
        std::memcpy(p + i, &x, std::min<std::size_t>(8, bytes - i));
    }
    return v;
}

// verify_sorted(&data_sorted, Some(&mut data_original)) (tests.rs:7-23)
template <class T>
void verify_sorted(const std::vector<T>& data, std::vector<T>& original) {
    std::sort(original.begin(), original.end());
    if (data.size() != original.size() || std::memcmp(data.data(), original.data(), data.size() * sizeof(T)) != 0) throw NotSorted();
}

// the 12 integer tests (tests.rs:25-131) are one shape
template <class T>
void integer_test(std::uint64_t seed) {
    std::vector<T> data_original = random_bits<T>(seed);
    std::vector<T> data_sorted = data_original;
    radix_sort::radix_sort(data_sorted);
    verify_sorted(data_sorted, data_original);
}

// f32::total_cmp / f64::total_cmp as the standard library defines them
template <class F, class I>
I total_key(F x) {
    I b;
    std::memcpy(&b, &x, sizeof b);
    using U = std::make_unsigned_t<I>;
    b ^= static_cast<I>(static_cast<U>(b >> (sizeof(I) * 8 - 1)) >> 1);
    return b;
}

// tests.rs:133-152 (f32) and :154-173 (f64)
template <class F, class I>
void float_test(std::uint64_t seed) {
    std::mt19937_64 rng(seed);
    std::uniform_real_distribution<F> unit(F(-1), F(1));  // Uniform::new(-1.0, 1.0)
    std::vector<F> data_original(N);
    for (auto& x : data_original) x = unit(rng);
    const F specials[5] = {F(0), -F(0), std::numeric_limits<F>::quiet_NaN(), std::numeric_limits<F>::infinity(),
                           -std::numeric_limits<F>::infinity()};
    for (F s : specials) data_original[rng() % N] = s;  // choose_mut(..) = special
    std::vector<F> data_sorted = data_original;
    radix_sort::radix_sort(data_sorted);
    std::sort(data_original.begin(), data_original.end(), [](F a, F b) { return total_key<F, I>(a) < total_key<F, I>(b); });
    for (std::size_t i = 0; i < N; ++i)
        if (total_key<F, I>(data_sorted[i]) != total_key<F, I>(data_original[i])) throw NotSorted();
}

// tests.rs:175-187: Vec<(u32, u32)>, stable by .0
void tuple_test(std::uint64_t seed) {
    std::mt19937_64 rng(seed);
    std::vector<std::pair<std::uint32_t, std::uint32_t>> data_original(N);
    for (auto& e : data_original) {
        const std::uint64_t x = rng();
        e = {static_cast<std::uint32_t>(x), static_cast<std::uint32_t>(x >> 32)};
    }
    // thread_rng gives ~116 equal keys in 1e6 draws; make sure some exist whatever the seed
    for (std::size_t i = 0; i < 1000; ++i) data_original[rng() % N].first = data_original[rng() % N].first;
    auto data_sorted = data_original;
    radix_sort::radix_sort(data_sorted);
    std::stable_sort(data_original.begin(), data_original.end(), [](const auto& a, const auto& b) { return a.first < b.first; });
    if (data_sorted != data_original) throw NotSorted();
}

struct Test {
    const char* name;
    std::function<void()> run;
};

}  // namespace

int main(int argc, char** argv) {
    const std::vector<Test> tests = {
        {"radix_sort_u8", [] { integer_test<std::uint8_t>(1); }},             // tests.rs:25-32
        {"radix_sort_u16", [] { integer_test<std::uint16_t>(2); }},           // :34-41
        {"radix_sort_u32", [] { integer_test<std::uint32_t>(3); }},           // :43-50
        {"radix_sort_u64", [] { integer_test<std::uint64_t>(4); }},           // :52-59
        {"radix_sort_u128", [] { integer_test<unsigned __int128>(5); }},      // :61-68
        {"radix_sort_usize", [] { integer_test<std::uint64_t>(6); }},         // :70-77
        {"radix_sort_i8", [] { integer_test<std::int8_t>(7); }},              // :79-86
        {"radix_sort_i16", [] { integer_test<std::int16_t>(8); }},            // :88-95
        {"radix_sort_i32", [] { integer_test<std::int32_t>(9); }},            // :97-104
        {"radix_sort_i64", [] { integer_test<std::int64_t>(10); }},           // :106-113
        {"radix_sort_i128", [] { integer_test<__int128>(11); }},              // :115-122
        {"radix_sort_isize", [] { integer_test<std::int64_t>(12); }},         // :124-131
        {"radix_sort_f32", [] { float_test<float, std::int32_t>(13); }},      // :133-152
        {"radix_sort_f64", [] { float_test<double, std::int64_t>(14); }},     // :154-173
        {"radix_sort_tuple", [] { tuple_test(15); }},                         // :175-187
    };
    const bool serial = argc > 1 && std::string(argv[1]) == "--test-threads=1";
    std::printf("running %zu tests\n", tests.size());
    std::atomic<int> failed{0};
    std::mutex out;
    auto one = [&](const Test& t) {
        std::string verdict = "ok";
        try {
            t.run();
        } catch (const std::exception& e) {
            verdict = std::string("FAILED (") + e.what() + ")";
            ++failed;
        }
        std::lock_guard<std::mutex> lk(out);
        std::printf("test %s ... %s\n", t.name, verdict.c_str());
    };
    if (serial) {
        for (const auto& t : tests) one(t);
    } else {
        std::vector<std::thread> th;
        for (const auto& t : tests) th.emplace_back(one, std::cref(t));
        for (auto& t : th) t.join();
    }
    std::printf("test result: %s. %zu passed; %d failed\n", failed ? "FAILED" : "ok", tests.size() - failed, failed.load());
    return failed;
}
