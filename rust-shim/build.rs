// build.rs — replaces the shaderc step of the reference's build.rs:17-73 for the compute path:
// compiles the sm_100a kernels with nvcc into libashcuda.a and links it (plus the CUDA runtime).
// UNBUILT here (no Rust toolchain in the image); flags mirror ash_renderer_b200/csrc/Makefile.
use std::{env, path::PathBuf, process::Command};

fn main() {
    let out = PathBuf::from(env::var("OUT_DIR").unwrap());
    let csrc = PathBuf::from(env::var("ASHCUDA_CSRC").unwrap_or_else(|_| "../ash_renderer_b200/csrc".into()));
    let nvcc = env::var("NVCC").unwrap_or_else(|_| "/usr/local/cuda/bin/nvcc".into());
    // the SRCS list of ash_renderer_b200/csrc/Makefile
    let sources = ["ashcuda.cu", "depth_pyramid.cu", "occlusion_cull.cu", "cull_compact_static.cu", "light_cull.cu", "post.cu", "post_f32.cu", "post_fast.cu", "post_pair.cu", "post_taa_tol.cu",
                   "post_up_pair.cu", "post_down_box.cu", "instancing.cu", "forward.cu", "compute_utils.cu", "view_graph.cu", "exact_taps.cu"];
    let mut objects = Vec::new();
    for s in sources {
        let obj = out.join(s).with_extension("o");
        let ok = Command::new(&nvcc)
            .args(["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-std=c++17", "-lineinfo", "-fmad=false",
                   "-prec-div=true", "-prec-sqrt=true", "-ftz=false", "-Xcompiler", "-fPIC", "-Xcompiler", "-ffp-contract=off", "-c"])
            .arg(csrc.join(s)).arg("-o").arg(&obj)
            .status().expect("nvcc not found").success();
        assert!(ok, "nvcc failed on {s}");
        println!("cargo:rerun-if-changed={}", csrc.join(s).display());
        objects.push(obj);
    }
    let lib = out.join("libashcuda.a");
    assert!(Command::new("ar").arg("crs").arg(&lib).args(&objects).status().unwrap().success());
    println!("cargo:rustc-link-search=native={}", out.display());
    println!("cargo:rustc-link-lib=static=ashcuda");
    println!("cargo:rustc-link-search=native=/usr/local/cuda/lib64");
    println!("cargo:rustc-link-lib=dylib=cudart");
    println!("cargo:rustc-link-lib=dylib=stdc++");
}
