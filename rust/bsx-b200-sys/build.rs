// build.rs of bsx-b200-sys.
//
// Default: link the prebuilt library.  BSX_B200_LIB_DIR names the directory holding libbsx_b200.so
// (0-kno-blobstreamx_b200/csrc after `make`); BSX_B200_INCLUDE_DIR the directory holding bsx_b200.h.
// Feature `build-cuda`: compile the kernels with nvcc for sm_100a only (`cc` with .cuda(true)) -- there is no other
// architecture and no CPU fallback.  Feature `bindgen`: regenerate the bindings from the header into OUT_DIR.
use std::env;
use std::path::PathBuf;

fn main() {
    let manifest = PathBuf::from(env::var("CARGO_MANIFEST_DIR").unwrap());
    let repo = manifest.join("..").join("..");
    let include = env::var("BSX_B200_INCLUDE_DIR").map(PathBuf::from).unwrap_or_else(|_| repo.join("include"));
    println!("cargo:rerun-if-env-changed=BSX_B200_LIB_DIR");
    println!("cargo:rerun-if-env-changed=BSX_B200_INCLUDE_DIR");
    println!("cargo:rerun-if-changed={}", include.join("bsx_b200.h").display());

    #[cfg(feature = "build-cuda")]
    {
        let csrc = repo.join("0-kno-blobstreamx_b200").join("csrc");
        let mut b = cc::Build::new();
        b.cuda(true)
            .cudart("shared")
            .flag("-std=c++17")
            .flag("-O3")
            .flag("-lineinfo")
            .flag("-gencode")
            .flag("arch=compute_100a,code=sm_100a")
            .include(&include);
        for f in ["api.cu", "api_batch.cu", "hash_kernels.cu", "ntt_kernels.cu", "fri_kernels.cu", "api_fri.cu",
                  "api_shard.cu", "eval_kernels.cu", "api_zs.cu", "api_proof.cu", "api_quotient.cu", "api_comm.cu", "api_noalloc.cu"] {
            println!("cargo:rerun-if-changed={}", csrc.join(f).display());
            b.file(csrc.join(f));
        }
        b.compile("bsx_b200");
        println!("cargo:rustc-link-lib=dylib=cudart");
    }
    #[cfg(not(feature = "build-cuda"))]
    {
        let lib_dir = env::var("BSX_B200_LIB_DIR").map(PathBuf::from)
            .unwrap_or_else(|_| repo.join("0-kno-blobstreamx_b200").join("csrc"));
        println!("cargo:rustc-link-search=native={}", lib_dir.display());
        println!("cargo:rustc-link-lib=dylib=bsx_b200");
        println!("cargo:rustc-link-arg=-Wl,-rpath,{}", lib_dir.display());
    }

    #[cfg(feature = "bindgen")]
    {
        let out = PathBuf::from(env::var("OUT_DIR").unwrap()).join("bindings.rs");
        bindgen::Builder::default()
            .header(include.join("bsx_b200.h").to_str().unwrap())
            .allowlist_function("bsx_.*")
            .allowlist_type("bsx_.*")
            .allowlist_var("BSX_.*")
            .generate()
            .expect("bindgen failed on bsx_b200.h")
            .write_to_file(out)
            .unwrap();
    }
}
