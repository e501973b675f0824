// Links the prebuilt libnorm_cuda.so (python build.py) or builds it with nvcc.
use std::{env, path::PathBuf, process::Command};

fn main() {
    if let Ok(dir) = env::var("NORM_CUDA_LIB_DIR") {
        println!("cargo:rustc-link-search=native={dir}");
    } else {
        let root = PathBuf::from(env::var("CARGO_MANIFEST_DIR").unwrap()).join("../..");
        let out = PathBuf::from(env::var("OUT_DIR").unwrap());
        let lib = out.join("libnorm_cuda.so");
        let status = Command::new(env::var("NVCC").unwrap_or_else(|_| "nvcc".into()))
            .args(["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-std=c++17", "-lineinfo", "-fmad=false"])
            .args(["-Xcompiler", "-fPIC,-O2,-ffp-contract=off", "-shared", "-cudart", "static", "-o"])
            .arg(&lib)
            .arg(root.join("norm_b200/csrc/norm_cuda.cu"))
            .status()
            .expect("nvcc not found");
        assert!(status.success(), "nvcc failed");
        println!("cargo:rustc-link-search=native={}", out.display());
    }
    println!("cargo:rustc-link-lib=dylib=norm_cuda");
    println!("cargo:rerun-if-env-changed=NORM_CUDA_LIB_DIR");
}
