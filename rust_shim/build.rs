// build.rs for the reference crate when linking the B200 renderer (NOT compiled in this repository).
fn main() {
    let dir = std::env::var("RRT_B200_LIB_DIR").expect("set RRT_B200_LIB_DIR to the directory holding librrt_b200.so");
    println!("cargo:rustc-link-search=native={dir}");
    println!("cargo:rustc-link-lib=dylib=rrt_b200");
    println!("cargo:rustc-link-arg=-Wl,-rpath,{dir}");
}
