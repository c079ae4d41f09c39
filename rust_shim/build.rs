// build.rs — what the reference's Rust crate (Rust/Cargo.toml) would add to link the B200 backend.
// Compiles the CUDA sources of this repo for sm_100a with nvcc and links them statically; no CPU
// fallback is built. SOURCE ONLY in this repository: the build image has no cargo/rustc, so this
// file has never been compiled here.
use std::env;
use std::path::PathBuf;
use std::process::Command;

fn main() {
    let out = PathBuf::from(env::var("OUT_DIR").unwrap());
    // RTW_B200_DIR points at a checkout of this repository
    let root = PathBuf::from(env::var("RTW_B200_DIR").expect("set RTW_B200_DIR to the b200 backend checkout"));
    let csrc = root.join("raytracing-in-a-weekend_b200/csrc");
    let nvcc = env::var("NVCC").unwrap_or_else(|_| "/usr/local/cuda/bin/nvcc".into());
    let common = ["-gencode", "arch=compute_100a,code=sm_100a", "-std=c++17", "-O3", "-lineinfo",
                  "-Xcompiler", "-fPIC,-ffp-contract=off,-fno-fast-math"];
    // rtw_kernels.cu three times: fast (FFMA contraction), strict (-fmad=false: rustc never fuses
    // a*b+c, so this build reproduces the crate's own f32 arithmetic), count (work counters)
    let variants: [(&str, &[&str]); 3] = [
        ("fast", &["-DRTW_VARIANT=fast"]),
        ("strict", &["-fmad=false", "-DRTW_VARIANT=strict", "-DRTW_STRICT_FP=1"]),
        ("count", &["-DRTW_VARIANT=count", "-DRTW_COUNT_WORK=1"]),
    ];
    let mut objs = Vec::new();
    for (name, extra) in variants.iter() {
        let o = out.join(format!("kernels_{name}.o"));
        run(Command::new(&nvcc).args(common).args(*extra).arg("-c").arg(csrc.join("rtw_kernels.cu")).arg("-o").arg(&o));
        objs.push(o);
    }
    // the rest of csrc/Makefile's OBJ list, one object per source (tests/test_rust_shim.py keeps the two lists equal):
    // rtw_lbvh.cu needs CUB, which ships with the CUDA toolkit (no extra include path)
    for (src, xcu) in [("rtw_api.cu", false), ("scene_build.cpp", true), ("rtw_lbvh.cu", false), ("rtw_post.cu", false),
                       ("rtw_multi.cu", false)] {
        let o = out.join(format!("{src}.o"));
        let mut c = Command::new(&nvcc);
        c.args(common);
        if xcu { c.args(["-x", "cu"]); }
        run(c.arg("-c").arg(csrc.join(src)).arg("-o").arg(&o));
        objs.push(o);
    }
    let lib = out.join("librtw_b200.a");
    run(Command::new("ar").arg("crs").arg(&lib).args(&objs));
    println!("cargo:rustc-link-search=native={}", out.display());
    println!("cargo:rustc-link-lib=static=rtw_b200");
    println!("cargo:rustc-link-search=native=/usr/local/cuda/lib64");
    println!("cargo:rustc-link-lib=dylib=cudart");
    println!("cargo:rustc-link-lib=dylib=stdc++");
    println!("cargo:rustc-link-lib=dylib=pthread");     // rtw_multi.cu: one host thread per device
    println!("cargo:rerun-if-changed={}", csrc.display());
}

fn run(c: &mut Command) {
    let st = c.status().expect("failed to spawn");
    assert!(st.success(), "command failed: {:?}", c);
}
