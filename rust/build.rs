// build.rs -- the nvcc step BASELINE.json's north_star asks for: compiles the sm_100a CUDA sources of
// ray-tracer-rust_b200/csrc into librtb200.so and links the crate against it.
// SOURCE ONLY: this image has no cargo/rustc, so this file is not exercised here; the tested build is
// ray-tracer-rust_b200/Makefile, which runs the same commands on the same file set: BOTH take every csrc/*.cu
// (tests/test_host.py::test_rust_build_uses_the_makefile_sources checks that neither hard-codes a list).
use std::env;
use std::path::PathBuf;
use std::process::Command;

fn main() {
    let root = PathBuf::from(env::var("CARGO_MANIFEST_DIR").unwrap());
    let pkg = root.join("..").join("ray-tracer-rust_b200");
    let out = PathBuf::from(env::var("OUT_DIR").unwrap());
    let nvcc = env::var("NVCC").unwrap_or_else(|_| "/usr/local/cuda/bin/nvcc".to_string());
    let arch = ["-gencode", "arch=compute_100a,code=sm_100a"];
    let mut objs = Vec::new();
    // every csrc/*.cu belongs to the library, exactly as `CU_SRCS := $(wildcard csrc/*.cu)` in the Makefile
    let mut names: Vec<String> = std::fs::read_dir(pkg.join("csrc")).expect("csrc/ missing")
        .filter_map(|e| e.ok()).map(|e| e.path())
        .filter(|p| p.extension().map_or(false, |x| x == "cu"))
        .map(|p| p.file_stem().unwrap().to_string_lossy().into_owned()).collect();
    names.sort();
    for name in names {
        let src = pkg.join("csrc").join(format!("{}.cu", name));
        let obj = out.join(format!("{}.o", name));
        println!("cargo:rerun-if-changed={}", src.display());
        // -fmad=false: the exact arithmetic of the reference must not be contracted (csrc/rtb_common.cuh)
        let ok = Command::new(&nvcc)
            .args(arch)
            .args(["-O3", "-std=c++17", "-lineinfo", "-fmad=false", "-Xcompiler", "-fPIC,-ffp-contract=off", "-c"])
            .arg(&src).arg("-o").arg(&obj)
            .status().expect("nvcc not found").success();
        assert!(ok, "nvcc failed on {}", src.display());
        objs.push(obj);
    }
    // every host/rtb_*.cpp (plain C++: OBJ readers, sample table, PNG writer), as `HOST_SRCS := $(wildcard host/rtb_*.cpp)` in the Makefile
    let mut hosts: Vec<std::path::PathBuf> = std::fs::read_dir(pkg.join("host")).expect("host/ missing")
        .filter_map(|e| e.ok()).map(|e| e.path())
        .filter(|p| p.extension().map_or(false, |x| x == "cpp") && p.file_name().unwrap().to_string_lossy().starts_with("rtb_")).collect();
    hosts.sort();
    for src in hosts {
        let obj = out.join(format!("{}.o", src.file_stem().unwrap().to_string_lossy()));
        let ok = Command::new("g++")
            .args(["-O2", "-std=c++17", "-fPIC", "-ffp-contract=off", "-pthread", "-c"])
            .arg(&src).arg("-o").arg(&obj)
            .status().expect("g++ not found").success();
        assert!(ok);
        objs.push(obj);
    }
    let ok = Command::new(&nvcc).args(arch).args(["-shared", "-cudart", "shared", "-o"])
        .arg(out.join("librtb200.so")).args(&objs)
        .status().unwrap().success();
    assert!(ok);
    println!("cargo:rustc-link-search=native={}", out.display());
    println!("cargo:rerun-if-changed={}", pkg.join("csrc").display());
    println!("cargo:rustc-link-lib=dylib=rtb200");
    println!("cargo:rustc-link-arg=-Wl,-rpath,{}", out.display());
}
