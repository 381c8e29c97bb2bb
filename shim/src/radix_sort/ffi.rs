//! `extern "C"` declarations of libb200sort.so -- one per symbol of include/b200sort.h that a
//! Rust caller of `RadixSort::radix_sort` (mod.rs:18-20,61) needs.  Checked against the header
//! by tests/test_shim.py (names, argument counts, argument classes).
//! UNVERIFIED: never compiled (no rustc where this was written; see shim/README.md).
#![allow(dead_code)]

use std::ffi::{c_char, c_int, c_uint, c_void, CStr};

/// `b200sort_kind` (include/b200sort.h): the built-in `RadixDigits` impls, radix_digits.rs:7-124.
#[repr(i32)]
#[derive(Clone, Copy, Debug, PartialEq, Eq)]
pub enum Kind {
    U8 = 0,
    U16 = 1,
    U32 = 2,
    U64 = 3,
    U128 = 4,
    I8 = 5,
    I16 = 6,
    I32 = 7,
    I64 = 8,
    I128 = 9,
    F32 = 10,
    F64 = 11,
}

/// Opaque communicator of the one-process multi-GPU entry points.
#[repr(C)]
pub struct Comm {
    _private: [u8; 0],
}

extern "C" {
    pub fn b200sort_version() -> c_int;
    pub fn b200sort_last_error() -> *const c_char;
    pub fn b200sort_key_bytes(kind: c_int) -> usize;
    pub fn b200sort_temp_bytes(n: usize, kind: c_int) -> usize;
    /// host slice in, host slice out: the drop-in for `slice.radix_sort()` on key-only slices
    pub fn b200sort_keys_host(keys: *mut c_void, n: usize, kind: c_int) -> c_int;
    /// device-resident keys, asynchronous on `stream`
    pub fn b200sort_keys_device(d_keys: *mut c_void, d_tmp: *mut c_void, tmp_bytes: usize, n: usize, kind: c_int, stream: *mut c_void) -> c_int;
    /// device-resident sort with the workspace inside the library image: for hosts like this one, which bind no
    /// CUDA runtime and so cannot allocate a `d_tmp` (limit: `b200sort_temp_bytes(n, kind)` <= 16 MiB)
    pub fn b200sort_keys_device_noalloc(d_keys: *mut c_void, n: usize, kind: c_int, stream: *mut c_void) -> c_int;
    /// device-resident sort whose temp buffer is taken and returned inside the call (library-owned stream-ordered pool)
    pub fn b200sort_keys_device_pooled(d_keys: *mut c_void, n: usize, kind: c_int, stream: *mut c_void) -> c_int;
    pub fn b200sort_keys_device_ex(d_keys: *mut c_void, d_tmp: *mut c_void, tmp_bytes: usize, n: usize, kind: c_int, stream: *mut c_void, flags: c_uint) -> c_int;
    /// `(K, V)` tuples (radix_digits.rs:126-136): whole records move, stable by the key at `key_off`
    pub fn b200sort_records_temp_bytes(n: usize, rec_bytes: usize, kind: c_int) -> usize;
    pub fn b200sort_records_host(recs: *mut c_void, n: usize, rec_bytes: usize, key_off: usize, kind: c_int) -> c_int;
    pub fn b200sort_records_device(d_recs: *mut c_void, d_tmp: *mut c_void, tmp_bytes: usize, n: usize, rec_bytes: usize, key_off: usize, kind: c_int, stream: *mut c_void) -> c_int;
    pub fn b200sort_set_host_pipeline_min_bytes(bytes: usize) -> c_int;
    pub fn b200sort_release_cached_memory() -> c_int;
    /// all GPUs of this process behind one call
    pub fn b200sort_mgpu_create(comm: *mut *mut Comm, ndev: c_int, devs: *const c_int) -> c_int;
    pub fn b200sort_mgpu_keys_device(comm: *mut Comm, d_shards: *const *const c_void, n_in: *const usize, kind: c_int, d_out: *mut *mut c_void, n_out: *mut usize) -> c_int;
    pub fn b200sort_mgpu_keys_host(comm: *mut Comm, keys: *mut c_void, n: usize, kind: c_int) -> c_int;
    pub fn b200sort_mgpu_destroy(comm: *mut Comm) -> c_int;
}

/// Non-zero status -> panic: the reference has no `Result`, failure is `expect` (mod.rs:68,106).
pub fn check(rc: c_int) {
    if rc != 0 {
        let msg = unsafe { CStr::from_ptr(b200sort_last_error()) }.to_string_lossy().into_owned();
        panic!("b200sort error {rc}: {msg}");
    }
}
