// Links the prebuilt CUDA library (built by `python -c 'import __graft_entry__ as g; g.build()'` into pathtrace_b200/lib/).
// NOT compiled on the build box of this repository (no cargo there): see rust/README.md.
use std::env;
use std::path::PathBuf;

fn main() {
    let dir = env::var("PATHTRACE_B200_LIB_DIR").map(PathBuf::from).unwrap_or_else(|_| {
        PathBuf::from(env::var("CARGO_MANIFEST_DIR").unwrap()).join("../../pathtrace_b200/lib")
    });
    println!("cargo:rustc-link-search=native={}", dir.display());
    println!("cargo:rustc-link-lib=dylib=pathtrace_b200");
    println!("cargo:rustc-link-arg=-Wl,-rpath,{}", dir.display());
    println!("cargo:rerun-if-env-changed=PATHTRACE_B200_LIB_DIR");
    println!("cargo:rerun-if-changed=../../include/pathtrace_b200.h");
}
