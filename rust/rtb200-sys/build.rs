// Links librtb200.so; RTB200_LIB_DIR = the directory that holds it (raytracing_rust_b200/ in the repo).
fn main() {
    println!("cargo:rerun-if-env-changed=RTB200_LIB_DIR");
    if let Ok(dir) = std::env::var("RTB200_LIB_DIR") {
        println!("cargo:rustc-link-search=native={}", dir);
        println!("cargo:rustc-link-arg=-Wl,-rpath,{}", dir);
    }
    println!("cargo:rustc-link-lib=dylib=rtb200");
}
