// UNBUILT here: this image has no rustc/cargo.  Kept in sync with include/rrt.h by
// tests/test_abi.py::test_rust_binding_declares_every_entry_point (names, field order, field types).
// build.rs
fn main() {
    let dir = std::env::var("RRT_LIB_DIR").expect("set RRT_LIB_DIR to the directory holding librrt.so");
    println!("cargo:rustc-link-search=native={dir}");
    println!("cargo:rustc-link-lib=dylib=rrt");
    println!("cargo:rustc-link-arg=-Wl,-rpath,{dir}");
}
