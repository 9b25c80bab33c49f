// Links libraysnail_b200.so.  RAYSNAIL_B200_LIB_DIR = the directory that holds it
// (<this repository>/raysnail_b200 after `make`).
fn main() {
    if let Ok(dir) = std::env::var("RAYSNAIL_B200_LIB_DIR") {
        println!("cargo:rustc-link-search=native={}", dir);
        println!("cargo:rustc-link-arg=-Wl,-rpath,{}", dir);
    }
    println!("cargo:rustc-link-lib=dylib=raysnail_b200");
    println!("cargo:rerun-if-env-changed=RAYSNAIL_B200_LIB_DIR");
}
