// Tells rustc where libnlc_commit.so lives: NLC_COMMIT_LIB_DIR, else <repo>/near-light-client_b200 next to this crate.
fn main() {
    let dir = std::env::var("NLC_COMMIT_LIB_DIR").unwrap_or_else(|_| {
        let here = std::path::PathBuf::from(std::env::var("CARGO_MANIFEST_DIR").unwrap());
        here.join("../../near-light-client_b200").to_string_lossy().into_owned()
    });
    println!("cargo:rustc-link-search=native={dir}");
    println!("cargo:rustc-link-lib=dylib=nlc_commit");
    println!("cargo:rustc-link-arg=-Wl,-rpath,{dir}");
    println!("cargo:rerun-if-env-changed=NLC_COMMIT_LIB_DIR");
}
