//! Safe wrapper replacing the reference's use of `tch` under `Member::predict`
//! (reference src/services.rs:455-456, 475-497, 513-524).
use b200_infer_sys as sys;
use std::ffi::{CStr, CString};
use std::ptr;

#[derive(Debug)]
pub struct Error { pub code: i32, pub message: String }

fn check(rc: i32) -> Result<(), Error> {
    if rc == sys::B2I_OK { return Ok(()); }
    let message = unsafe { CStr::from_ptr(sys::b2i_last_error()) }.to_string_lossy().into_owned();
    Err(Error { code: rc, message })
}

/// A set of GPUs on this member (replaces `tch::Device::Cpu`).
pub struct Ctx(*mut sys::b2i_ctx);
unsafe impl Send for Ctx {}
unsafe impl Sync for Ctx {}

impl Ctx {
    pub fn all_gpus() -> Result<Self, Error> {
        let ids: Vec<i32> = (0..unsafe { sys::b2i_device_count() }).collect();
        Self::new(&ids)
    }
    pub fn new(device_ids: &[i32]) -> Result<Self, Error> {
        let mut p = ptr::null_mut();
        check(unsafe { sys::b2i_init(device_ids.as_ptr(), device_ids.len() as i32, &mut p) })?;
        Ok(Ctx(p))
    }
    /// First half of `imagenet::load_image_and_resize` after the host-side JPEG decode.
    pub fn resize_crop(&self, rgb: &[u8], h: u32, w: u32) -> Result<Vec<u8>, Error> {
        assert_eq!(rgb.len(), (h * w * 3) as usize);
        let mut out = vec![0u8; sys::B2I_IMAGE_BYTES];
        check(unsafe { sys::b2i_resize_crop_u8(self.0, 0, rgb.as_ptr(), h as i32, w as i32, out.as_mut_ptr()) })?;
        Ok(out)
    }
}
impl Drop for Ctx { fn drop(&mut self) { unsafe { sys::b2i_shutdown(self.0) } } }

/// One loaded network (replaces `Mutex<Box<dyn ModuleT>>`; calls are serialised inside the library).
pub struct Model(*mut sys::b2i_model);
unsafe impl Send for Model {}
unsafe impl Sync for Model {}

impl Model {
    /// `resnet::resnet18(..)` / `alexnet::alexnet(..)` + `vs.load(path)`.
    pub fn load(ctx: &Ctx, name: &str, ot_path: &str, max_batch: i32) -> Result<Self, Error> {
        let mut cfg = unsafe { std::mem::zeroed::<sys::b2i_model_config>() };
        unsafe { sys::b2i_default_config(&mut cfg) };
        cfg.max_batch = max_batch;
        let (n, p) = (CString::new(name).unwrap(), CString::new(ot_path).unwrap());
        let mut m = ptr::null_mut();
        check(unsafe { sys::b2i_model_load(ctx.0, n.as_ptr(), p.as_ptr(), &cfg, &mut m) })?;
        Ok(Model(m))
    }
    /// `forward_t(.., false).softmax(-1)` + `imagenet::top(.., k)` over a batch of 224x224x3 crops.
    /// `devices`: None = all GPUs, Some((first, count)) = the job's share (services.rs:199-211).
    pub fn predict(&self, crops: &[u8], k: usize, devices: Option<(i32, i32)>) -> Result<Vec<Vec<(f64, usize)>>, Error> {
        let n = crops.len() / sys::B2I_IMAGE_BYTES;
        let (mut idx, mut prob) = (vec![0i32; n * k], vec![0f32; n * k]);
        let rc = unsafe {
            match devices {
                None => sys::b2i_predict_u8(self.0, crops.as_ptr(), n as i32, k as i32, idx.as_mut_ptr(), prob.as_mut_ptr(), ptr::null_mut()),
                Some((f, c)) => sys::b2i_predict_u8_on(self.0, f, c, crops.as_ptr(), n as i32, k as i32, idx.as_mut_ptr(), prob.as_mut_ptr(), ptr::null_mut()),
            }
        };
        check(rc)?;
        Ok((0..n).map(|i| (0..k).map(|j| (prob[i * k + j] as f64, idx[i * k + j] as usize)).collect()).collect())
    }
    /// The body of `Member::predict` from the file contents on (services.rs:484-495): encoded JPEG streams in,
    /// one `Some(top-k)` per stream that decoded, `None` for one that did not (the caller maps any `None`
    /// to the RPC-level `None` of `.ok()?` at :492, or keeps per-id accounting in a batched `run_job`).
    pub fn predict_jpeg(&self, blobs: &[&[u8]], k: usize, devices: Option<(i32, i32)>) -> Result<Vec<Option<Vec<(f64, usize)>>>, Error> {
        let n = blobs.len();
        let ptrs: Vec<*const u8> = blobs.iter().map(|b| b.as_ptr()).collect();
        let lens: Vec<usize> = blobs.iter().map(|b| b.len()).collect();
        let (mut idx, mut prob, mut status) = (vec![0i32; n * k], vec![0f32; n * k], vec![0i32; n]);
        let (f, c) = devices.unwrap_or((0, 0));
        check(unsafe {
            sys::b2i_predict_jpeg(self.0, f, c, ptrs.as_ptr(), lens.as_ptr(), n as i32, k as i32, idx.as_mut_ptr(),
                                  prob.as_mut_ptr(), ptr::null_mut(), status.as_mut_ptr())
        })?;
        Ok((0..n).map(|i| if status[i] != sys::B2I_OK { None } else {
            Some((0..k).map(|j| (prob[i * k + j] as f64, idx[i * k + j] as usize)).collect())
        }).collect())
    }
    /// After `train` has copied a new `<model>.ot` to this member (main.rs:231-256): swap the weights in.
    pub fn reload(&self, ot_path: &str) -> Result<(), Error> {
        let p = CString::new(ot_path).unwrap();
        check(unsafe { sys::b2i_model_reload(self.0, p.as_ptr()) })
    }
}
impl Drop for Model { fn drop(&mut self) { unsafe { sys::b2i_model_free(self.0) } } }
