// build.rs -- links libb200sort.so when the crate is built with `--features b200`.
// Hand-written declarations (src/radix_sort/ffi.rs): bindgen is not needed.
// UNVERIFIED: never run through cargo (see shim/README.md).
fn main() {
    println!("cargo:rerun-if-env-changed=B200SORT_LIB_DIR");
    if std::env::var_os("CARGO_FEATURE_B200").is_none() {
        return;
    }
    let dir = std::env::var("B200SORT_LIB_DIR")
        .expect("feature `b200`: set B200SORT_LIB_DIR to the directory that holds libb200sort.so");
    println!("cargo:rustc-link-search=native={dir}");
    println!("cargo:rustc-link-lib=dylib=b200sort");
    // the test binaries and `cargo run` find the library without LD_LIBRARY_PATH
    println!("cargo:rustc-link-arg=-Wl,-rpath,{dir}");
}
