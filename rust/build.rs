// build.rs — links librt_b200.so (built by gpu_ray_tracing_b200/csrc/build.sh of the B200 repository).
// RT_B200_LIB_DIR names the directory that holds the library; it is also put on the run-time search path so that
// `cargo run` finds it without LD_LIBRARY_PATH.
fn main() {
    println!("cargo:rerun-if-env-changed=RT_B200_LIB_DIR");
    let dir = std::env::var("RT_B200_LIB_DIR")
        .expect("set RT_B200_LIB_DIR to the directory that contains librt_b200.so");
    println!("cargo:rustc-link-search=native={dir}");
    println!("cargo:rustc-link-lib=dylib=rt_b200");
    println!("cargo:rustc-link-arg=-Wl,-rpath,{dir}");
}
