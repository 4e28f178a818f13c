// never compiled here. Builds arrow-ballista_b200/libbgpu.so (nvcc -gencode arch=compute_100a,code=sm_100a, see
// arrow-ballista_b200/csrc/Makefile) and links it; generates the codec's messages from proto/bgpu.proto.
fn main() {
    let dir = std::env::var("BGPU_DIR").unwrap_or_else(|_| "../arrow-ballista_b200".to_string());
    let ok = std::process::Command::new("make").arg("-j8").arg("-C").arg(format!("{dir}/csrc")).status()
        .expect("make not found").success();
    assert!(ok, "building libbgpu.so failed");
    println!("cargo:rustc-link-search=native={dir}");
    println!("cargo:rustc-link-lib=dylib=bgpu");
    println!("cargo:rerun-if-changed={dir}/csrc");
    prost_build::compile_protos(&[format!("{dir}/proto/bgpu.proto")], &[format!("{dir}/proto")]).unwrap();
}
