// build_rcull.rs -- the step a maintainer adds to farnoy/renderer's build.rs (next to the glslc step,
// build.rs:29-76): compile librcull.so with nvcc for sm_100a and link it.  Call build_rcull() from main().
// Source only: never compiled on the box that wrote it (no rustc there).  `rcull/` is this repository
// checked out (or vendored) at the root of the renderer tree.
pub fn build_rcull() {
    let out = std::path::PathBuf::from(std::env::var("OUT_DIR").unwrap());
    let src = "rcull/renderer_b200/csrc";
    let status = std::process::Command::new("nvcc")
        .args(["-std=c++17", "-O3", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo",
               "-fmad=false", "-Xcompiler", "-fPIC,-ffp-contract=off", "--cudart", "static",
               "-Ircull/include", "-shared", "-o"])
        .arg(out.join("librcull.so"))
        .args([format!("{src}/rcull.cu"), format!("{src}/rcull_host.cpp")])
        .status()
        .expect("nvcc not found");
    assert!(status.success(), "nvcc failed");
    println!("cargo:rustc-link-search=native={}", out.display());
    println!("cargo:rustc-link-lib=dylib=rcull");
    println!("cargo:rerun-if-changed={src}");
    println!("cargo:rerun-if-changed=rcull/include/rcull.h");
}
