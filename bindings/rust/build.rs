// build.rs — compiles the CUDA sources with nvcc and links the result.
// SFP_ROOT points at a checkout of this repository (default: two directories up).
use std::{env, path::PathBuf, process::Command};

fn main() {
    let out = PathBuf::from(env::var("OUT_DIR").unwrap());
    let root = env::var("SFP_ROOT").map(PathBuf::from).unwrap_or_else(|_| {
        PathBuf::from(env::var("CARGO_MANIFEST_DIR").unwrap()).join("../..")
    });
    let csrc = root.join("starframe_b200/csrc");
    let inc = root.join("include");
    let lib = out.join("libsfp_b200.so");
    let status = Command::new(env::var("NVCC").unwrap_or_else(|_| "nvcc".into()))
        .args([
            "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
            // not optional: bit-identical results with the CPU oracle depend on it
            "--fmad=false",
            "-Xcompiler", "-fPIC,-ffp-contract=off,-fvisibility=hidden",
            "--expt-relaxed-constexpr", "-shared",
        ])
        .arg("-I").arg(&inc)
        .arg("-I").arg(&csrc)
        .arg("-o").arg(&lib)
        .arg(csrc.join("sfp_world.cu"))
        .arg(csrc.join("sfp_batch.cu"))
        .status()
        .expect("nvcc not found (set NVCC)");
    assert!(status.success(), "nvcc failed");
    println!("cargo:rustc-link-search=native={}", out.display());
    println!("cargo:rustc-link-lib=dylib=sfp_b200");
    for f in ["sfp_world.cu", "sfp_batch.cu", "sfp_device.cuh", "sfp_sort.cuh"] {
        println!("cargo:rerun-if-changed={}", csrc.join(f).display());
    }
    for f in ["sfp.h", "sfp_math.h", "sfp_host.h"] {
        println!("cargo:rerun-if-changed={}", inc.join(f).display());
    }
}
