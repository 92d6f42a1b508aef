// Links libb200infer.so built by distributed-machine-learning-cluster_b200/build.sh.
fn main() {
    let dir = std::env::var("B200_INFER_LIB_DIR")
        .unwrap_or_else(|_| "../../../distributed-machine-learning-cluster_b200".to_string());
    println!("cargo:rustc-link-search=native={}", dir);
    println!("cargo:rustc-link-lib=dylib=b200infer");
    println!("cargo:rerun-if-env-changed=B200_INFER_LIB_DIR");
}
