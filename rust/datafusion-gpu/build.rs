// libdfgpu.so is built by `python -c 'import __graft_entry__ as g; g.build()'` (nvcc, sm_100a) into
// datafusion_upstream_b200/; DFGPU_LIB_DIR overrides where the linker looks for it.
fn main() {
    let dir = std::env::var("DFGPU_LIB_DIR").unwrap_or_else(|_| "../../datafusion_upstream_b200".to_string());
    println!("cargo:rustc-link-search=native={dir}");
    println!("cargo:rustc-link-lib=dylib=dfgpu");
    println!("cargo:rerun-if-env-changed=DFGPU_LIB_DIR");
}
