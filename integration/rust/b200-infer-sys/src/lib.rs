//! Raw declarations of include/b200_infer.h (one per entry point the reference path needs).
#![allow(non_camel_case_types)]
use std::os::raw::{c_char, c_double, c_float, c_int, c_longlong, c_void};

#[repr(C)] pub struct b2i_ctx { _private: [u8; 0] }
#[repr(C)] pub struct b2i_model { _private: [u8; 0] }

#[repr(C)]
#[derive(Clone, Copy, Debug)]
pub struct b2i_model_config {
    pub precision: c_int,
    pub max_batch: c_int,
    pub micro_batch: c_int,
    pub alexnet_geometry: c_int,
    pub conv_mode: c_int,
    pub weights_tma: c_int,
    pub tile_n_max: c_int,
    pub keep_tensors: c_int,
    pub fused_stem: c_int,
    pub patch_conv: c_int,
    pub pair_mma: c_int,
    pub fuse_downsample: c_int,
}

pub const B2I_OK: c_int = 0;
pub const B2I_EINVAL: c_int = -1;
pub const B2I_EIO: c_int = -2;
pub const B2I_ECUDA: c_int = -3;
pub const B2I_ENODEV: c_int = -4;
pub const B2I_ENOMEM: c_int = -5;
pub const B2I_EINTERNAL: c_int = -6;
pub const B2I_EDECODE: c_int = -7;
pub const B2I_PREC_BF16: c_int = 0;
pub const B2I_PREC_FP32: c_int = 1;
pub const B2I_PREC_FP16: c_int = 2;
pub const B2I_IMAGE_BYTES: usize = 224 * 224 * 3;
pub const B2I_CLASSES: usize = 1000;

extern "C" {
    pub fn b2i_last_error() -> *const c_char;
    pub fn b2i_device_count() -> c_int;
    pub fn b2i_init(device_ids: *const c_int, n_devices: c_int, out: *mut *mut b2i_ctx) -> c_int;
    pub fn b2i_shutdown(ctx: *mut b2i_ctx);
    pub fn b2i_default_config(cfg: *mut b2i_model_config);
    pub fn b2i_model_load(ctx: *mut b2i_ctx, model_name: *const c_char, ot_path: *const c_char,
                          cfg: *const b2i_model_config, out: *mut *mut b2i_model) -> c_int;
    pub fn b2i_model_free(m: *mut b2i_model);
    pub fn b2i_model_reload(m: *mut b2i_model, ot_path: *const c_char) -> c_int;
    pub fn b2i_model_flops_per_image(m: *const b2i_model) -> c_double;
    pub fn b2i_predict_u8(m: *mut b2i_model, images: *const u8, n: c_int, k: c_int,
                          idx: *mut i32, prob: *mut c_float, logits: *mut c_float) -> c_int;
    pub fn b2i_predict_u8_on(m: *mut b2i_model, first_dev: c_int, n_dev: c_int, images: *const u8,
                             n: c_int, k: c_int, idx: *mut i32, prob: *mut c_float,
                             logits: *mut c_float) -> c_int;
    pub fn b2i_predict_raw_u8(m: *mut b2i_model, first_dev: c_int, n_dev: c_int, images: *const *const u8,
                              heights: *const c_int, widths: *const c_int, n: c_int, k: c_int,
                              idx: *mut i32, prob: *mut c_float, logits: *mut c_float) -> c_int;
    pub fn b2i_predict_raw_u8_status(m: *mut b2i_model, first_dev: c_int, n_dev: c_int, images: *const *const u8,
                                     heights: *const c_int, widths: *const c_int, n: c_int, k: c_int,
                                     idx: *mut i32, prob: *mut c_float, logits: *mut c_float,
                                     status: *mut c_int) -> c_int;
    pub fn b2i_predict_jpeg(m: *mut b2i_model, first_dev: c_int, n_dev: c_int, jpeg: *const *const u8,
                            len: *const usize, n: c_int, k: c_int, idx: *mut i32, prob: *mut c_float,
                            logits: *mut c_float, status: *mut c_int) -> c_int;
    pub fn b2i_jpeg_decode_u8(ctx: *mut b2i_ctx, dev: c_int, jpeg: *const u8, len: usize, out: *mut u8,
                              cap: usize, h: *mut c_int, w: *mut c_int) -> c_int;
    pub fn b2i_submit_u8(m: *mut b2i_model, images: *const u8, n: c_int, k: c_int, idx: *mut i32,
                         prob: *mut c_float, logits: *mut c_float, ticket: *mut c_longlong) -> c_int;
    pub fn b2i_wait(m: *mut b2i_model, ticket: c_longlong) -> c_int;
    pub fn b2i_resize_crop_u8(ctx: *mut b2i_ctx, dev: c_int, image: *const u8, h: c_int, w: c_int,
                              out224: *mut u8) -> c_int;
    pub fn b2i_resize_crop_u8_static(ctx: *mut b2i_ctx, dev: c_int, d_image: *const c_void, h: c_int,
                                     w: c_int, d_out224: *mut c_void, cuda_stream: *mut c_void) -> c_int;
    pub fn b2i_host_alloc(bytes: usize) -> *mut c_void;
    pub fn b2i_host_free(p: *mut c_void);
}
