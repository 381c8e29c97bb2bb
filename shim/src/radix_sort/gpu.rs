//! GPU body of `RadixSort::radix_sort` (replaces mod.rs:61-176 when the crate is built with
//! `--features b200`): one call into libb200sort.so per slice.
//! UNVERIFIED: never compiled (no rustc where this was written; see shim/README.md).
//!
//! `impl<T: RadixDigits> RadixSort<T> for [T]` (mod.rs:61) is generic over user `get_digit`
//! code, which a GPU cannot run.  `GpuKey` names, for every built-in `RadixDigits` impl
//! (radix_digits.rs:7-136), the key kind of the C ABI and where the key sits in the element;
//! the GPU impl is bounded by it, so a user-defined impl is rejected at compile time (the C
//! ABI's B200SORT_ERR_UNSUPPORTED_KIND).

use super::ffi::{self, Kind};
use super::{RadixDigits, RadixSort};

pub trait GpuKey: RadixDigits {
    /// key kind of the element's key field
    const KIND: Kind;
    /// byte offset of the key inside the element (0 for scalars)
    const KEY_OFFSET: usize = 0;
    /// true for `(K, V)`: the whole element moves, the sort is stable by the key
    const IS_RECORD: bool = false;
}

macro_rules! gpu_key {
    ($($t:ty => $k:ident),* $(,)?) => { $( impl GpuKey for $t { const KIND: Kind = Kind::$k; } )* };
}
// radix_digits.rs:7-45 unsigned, :55-93 signed, :103-124 floats
gpu_key!(u8 => U8, u16 => U16, u32 => U32, u64 => U64, u128 => U128,
         i8 => I8, i16 => I16, i32 => I32, i64 => I64, i128 => I128,
         f32 => F32, f64 => F64);
// radix_digits.rs:47-53, 95-101: NUMBER_OF_DIGITS = size_of::<usize>()
#[cfg(target_pointer_width = "64")]
gpu_key!(usize => U64, isize => I64);
#[cfg(target_pointer_width = "32")]
gpu_key!(usize => U32, isize => I32);

// radix_digits.rs:126-136: the digit of a tuple is the digit of `.0`; `.1` rides along.
// Tuples are repr(Rust): the offset of `.0` is computed, never assumed.
// Bounds spelled out as in the reference's own tuple impl (radix_digits.rs:126-130: `T: RadixDigits + Send + Sync`,
// `U: Send + Sync`), which is what the supertrait `RadixDigits` of `GpuKey` requires of `(T, U)`.  (`RadixDigits:
// Send + Sync` at radix_digits.rs:1 already implies them for T; repeating them keeps the two impls side by side.)
// The library sorts two record shapes (b200sort_api.cu, check_record_shape): a 4-byte key in an 8-byte
// record and an 8-byte key in a 16-byte record.  Any other `(T, U)` is refused HERE, when the constant is
// evaluated for the concrete types (a compile error at the call site), instead of a panic at run time
// where the reference would have sorted.
impl<T: GpuKey + Send + Sync, U: Send + Sync> GpuKey for (T, U) {
    const KIND: Kind = T::KIND;
    const KEY_OFFSET: usize = {
        let rec = std::mem::size_of::<(T, U)>();
        let key = std::mem::size_of::<T>();
        assert!(
            (rec == 8 && key == 4) || (rec == 16 && key == 8),
            "libb200sort sorts (K, V) records of 8 bytes with a 4-byte key or of 16 bytes with an 8-byte key only"
        );
        assert!(!T::IS_RECORD, "nested tuples are not a record shape of the library");
        std::mem::offset_of!((T, U), 0) + T::KEY_OFFSET
    };
    const IS_RECORD: bool = true;
}

impl<T: GpuKey> RadixSort<T> for [T] {
    fn radix_sort(&mut self) {
        let n = self.len();
        let p = self.as_mut_ptr().cast::<std::ffi::c_void>();
        // (n == 0 returns at once in the library; the reference panics there, mod.rs:92)
        let rc = unsafe {
            if T::IS_RECORD {
                ffi::b200sort_records_host(p, n, std::mem::size_of::<T>(), T::KEY_OFFSET, T::KIND as i32)
            } else {
                ffi::b200sort_keys_host(p, n, T::KIND as i32)
            }
        };
        ffi::check(rc);
    }
}
