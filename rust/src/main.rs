//! The reference's `main` (src/main.rs:319-362) with `render(scene)` replaced by the C ABI of librtb200.so.
//! SOURCE ONLY (no Rust toolchain in the build image); the tested twin is ray-tracer-rust_b200/host/rt_main.cpp.
mod ffi;
use std::env;
use std::ffi::{CStr, CString};
use std::ptr;

const NB_RAY: u32 = 1;                 // src/main.rs:38
const NB_LIGHT_SAMPLE: u32 = 100;      // src/main.rs:39
const NB_RANDOM_SAMPLE: u32 = 2_000_000; // src/main.rs:40

fn check(rc: i32) {
    if rc != ffi::RTB_OK {
        let msg = unsafe { CStr::from_ptr(ffi::rtb_last_error()) };
        panic!("rtb200 error {}: {}", rc, msg.to_string_lossy());
    }
}

fn main() {
    println!("Building scene");                                                    // :322
    let (mut verts, mut colors): (Vec<f32>, Vec<f32>) = (Vec::new(), Vec::new());
    // every argument is an OBJ path (src/main.rs:328-331), except the optional `--gpus N` (an extension: the reference's
    // render() sizes its own worker pool, src/main.rs:269; here N devices of the box share the frame's row bands)
    let mut n_gpus: u32 = 1;
    let mut paths: Vec<String> = Vec::new();
    let mut args = env::args().skip(1);
    while let Some(a) = args.next() {
        if a == "--gpus" { n_gpus = args.next().and_then(|v| v.parse().ok()).unwrap_or(1).max(1); } else { paths.push(a); }
    }
    for path in paths {                                                            // :328-331
        let c = CString::new(path).unwrap();
        let (mut p, mut n) = (ptr::null_mut::<f32>(), 0u32);
        let rc = unsafe { ffi::rtb_import_obj(c.as_ptr(), &mut p, &mut n) };
        if rc != ffi::RTB_OK && n == 0 {                                           // :115-121 prints and continues
            println!("{}", unsafe { CStr::from_ptr(ffi::rtb_last_error()) }.to_string_lossy());
            continue;
        }
        if n > 0 && !p.is_null() {                                                 // an OBJ without faces: RTB_OK, n = 0, p = NULL
            verts.extend_from_slice(unsafe { std::slice::from_raw_parts(p, n as usize * 9) });
            colors.extend(std::iter::repeat(1.0f32).take(n as usize * 3));         // :143
        }
        unsafe { ffi::rtb_free(p as *mut _) };
    }
    verts.extend_from_slice(&[-10000.0, 0.0, -10000.0, 10000.0, 0.0, -10000.0, 0.0, 0.0, 10000.0]);   // :102-111
    colors.extend_from_slice(&[0.5, 0.5, 0.5]);

    let mut samples = vec![0f32; NB_RANDOM_SAMPLE as usize * 2];                   // :253,260-265
    unsafe { ffi::rtb_fill_samples(0x5EED0001, NB_RANDOM_SAMPLE, samples.as_mut_ptr()) };
    let mut scenes: Vec<*mut ffi::RtbScene> = Vec::new();                          // one replica of the scene per GPU
    for device in 0..n_gpus as i32 {
        let desc = ffi::RtbSceneDesc {
            n_tris: (verts.len() / 9) as u32, verts: verts.as_ptr(), colors: colors.as_ptr(), tie_rank: ptr::null(),
            light: [-10.0, 300.0, -10.0, 10.0, 300.0, -10.0, 0.0, 300.0, 0.0],     // :337-343
            device,
        };
        let mut sc = ptr::null_mut();
        check(unsafe { ffi::rtb_scene_create(&desc, &mut sc) });                   // replaces BVH::new, :357
        scenes.push(sc);
    }
    check(unsafe { ffi::rtb_scenes_set_samples(scenes.as_ptr(), n_gpus, samples.as_ptr(), NB_RANDOM_SAMPLE) });
    let scene = scenes[0];

    let (width, height) = (1920u32, 1080u32);                                      // :350-351
    let params = ffi::RtbRenderParams {
        width, height, nb_ray: NB_RAY, nb_light_sample: NB_LIGHT_SAMPLE,
        eye: [0.0, 100.0, 200.0], look_at: [0.0, 0.0, -100000.0], up: [0.0, 1.0, 0.0], distance: 288.0,   // :353-356
        band_rows: 4, rank: 0, world: 1, flags: 0,
    };
    println!("Rendering...");                                                      // :360
    let mut img = vec![0u8; (width * height * 3) as usize];
    let mut stats = ffi::RtbStats::default();
    let t0 = std::time::Instant::now();
    if n_gpus > 1 {                                                                // render() on N GPUs: row bands b % N, joined inside the call
        check(unsafe { ffi::rtb_render_multi(scenes.as_ptr(), n_gpus, &params, img.as_mut_ptr(), &mut stats) });
    } else {
        check(unsafe { ffi::rtb_render(scene, &params, img.as_mut_ptr(), &mut stats) });   // replaces render(), :361
    }
    let secs = t0.elapsed().as_secs_f64();
    println!("Rendered in {} seconds", secs);                                      // :305-306
    println!("Throughput {}M ray/s", stats.primary_rays as f64 / 1e6 / secs);      // :308-309 (primary rays)
    println!("Writting image to disk");                                            // :312 [sic]
    let out = CString::new("output.png").unwrap();
    check(unsafe { ffi::rtb_write_png(out.as_ptr(), img.as_ptr(), width, height) });   // :313-315
    for sc in scenes { unsafe { ffi::rtb_scene_destroy(sc) }; }
}
