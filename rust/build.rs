// build.rs — compiles the CUDA sources of the XPBD step path with nvcc for sm_100a and links them into the crate.
// Drop this file next to Cargo.toml of jim-ec/constraint_solver together with `csrc/` and `include/` from this
// repository.  No PyTorch, no Triton, no wgpu compute: plain nvcc + the CUDA runtime.
use std::{env, path::PathBuf, process::Command};

fn main() {
    let out = PathBuf::from(env::var("OUT_DIR").unwrap());
    let csrc = PathBuf::from("csrc");
    let nvcc = env::var("NVCC").unwrap_or_else(|_| "nvcc".into());
    let mut objects = Vec::new();
    for entry in std::fs::read_dir(&csrc).expect("csrc/ missing") {
        let path = entry.unwrap().path();
        let ext = path.extension().and_then(|e| e.to_str()).unwrap_or("");
        if ext != "cu" && ext != "cpp" {
            continue;
        }
        println!("cargo:rerun-if-changed={}", path.display());
        let obj = out.join(path.file_stem().unwrap()).with_extension("o");
        let status = Command::new(&nvcc)
            .args([
                "-gencode", "arch=compute_100a,code=sm_100a",   // B200 only
                "-lineinfo", "-O3", "-std=c++17",
                "-fmad=false",                                    // bit parity with the fp32 CPU solver
                "-Xcompiler", "-fPIC,-ffp-contract=off",
                "-Iinclude", "-c", "-o",
            ])
            .arg(&obj)
            .arg(&path)
            .status()
            .expect("failed to run nvcc");
        assert!(status.success(), "nvcc failed on {}", path.display());
        objects.push(obj);
    }
    let lib = out.join("libxpbd_b200.a");
    let status = Command::new("ar").arg("crs").arg(&lib).args(&objects).status().expect("failed to run ar");
    assert!(status.success());
    let cuda = env::var("CUDA_HOME").unwrap_or_else(|_| "/usr/local/cuda".into());
    println!("cargo:rustc-link-search=native={}", out.display());
    println!("cargo:rustc-link-search=native={}/lib64", cuda);
    println!("cargo:rustc-link-lib=static=xpbd_b200");
    println!("cargo:rustc-link-lib=dylib=cudart");
    println!("cargo:rustc-link-lib=dylib=stdc++");
    println!("cargo:rerun-if-changed=include/xpbd_b200.h");
}
