// renderer.rs — safe wrapper over the raw FFI of rt_b200.rs.
//
// `Renderer` owns what the render world of the reference owned for this pass (src/lib.rs:138-229): the accumulation
// image, the sphere storage buffer + count and the camera uniform, all device side behind one opaque handle.  Host
// memory is only ever borrowed for the duration of a call, as with wgpu's `create_buffer_with_data` (src/lib.rs:158-194).
// Errors come back as negative status codes + rt_last_error(); nothing aborts across the boundary, so the caller decides
// whether to panic (as the reference does, src/lib.rs:216-217, 356-358) or to map to NodeRunError.
//
// Not compiled in the repository's build image (no Rust toolchain there).
use crate::camera::SceneCamera;
use crate::rt_b200 as ffi;
use crate::scene::sphere::GpuSphere;
use std::ffi::CStr;
use std::ptr;

#[derive(Debug, Clone)]
pub struct RtError {
    pub code: i32,
    pub message: String,
}

impl std::fmt::Display for RtError {
    fn fmt(&self, f: &mut std::fmt::Formatter<'_>) -> std::fmt::Result {
        write!(f, "rt_b200 error {}: {}", self.code, self.message)
    }
}
impl std::error::Error for RtError {}

fn check(code: i32) -> Result<(), RtError> {
    if code == ffi::RT_OK {
        return Ok(());
    }
    // thread-local, never null (include/rt_abi.h)
    let message = unsafe { CStr::from_ptr(ffi::rt_last_error()) }.to_string_lossy().into_owned();
    Err(RtError { code, message })
}

/// Arithmetic of the per-ray O(1) work (rt_set_precision).
#[derive(Clone, Copy, Debug, PartialEq, Eq)]
pub enum Precision {
    /// one IEEE rounding per WGSL operator: bit-identical to the repository's CPU oracle (default)
    Exact,
    /// hardware approximations + FMA contraction in scatter / sky colour / running mean; primary rays exact
    FastShading,
    /// ray generation as well (silhouette samples may flip)
    FastAll,
}

pub struct Renderer {
    raw: *mut ffi::RtRenderer,
    pub width: u32,
    pub height: u32,
    /// the row window this renderer owns ([row_begin, row_begin + row_count) of the full image): a multi-GPU shard
    pub row_begin: u32,
    pub row_count: u32,
}

// The handle has no thread affinity (every entry sets its device first); one caller at a time, which &mut / Bevy's
// resource borrowing gives.  The render sub-app may live on another thread than the main world (pipelined rendering).
unsafe impl Send for Renderer {}
unsafe impl Sync for Renderer {}

impl Renderer {
    pub fn new(device: i32, width: u32, height: u32) -> Result<Self, RtError> {
        Self::window(device, width, height, 0, height)
    }

    pub fn window(device: i32, width: u32, height: u32, row_begin: u32, row_count: u32) -> Result<Self, RtError> {
        let mut raw = ptr::null_mut();
        check(unsafe { ffi::rt_create_window(device, width, height, row_begin, row_count, &mut raw) })?;
        Ok(Self { raw, width, height, row_begin, row_count })
    }

    /// prepare_sphere_buffer (src/lib.rs:177-207): the padded Vec and the live count
    pub fn set_scene(&self, spheres: &[GpuSphere], count: u32) -> Result<(), RtError> {
        check(unsafe { ffi::rt_set_scene(self.raw, spheres.as_ptr(), spheres.len() as u32, count) })
    }

    /// the `init` dispatch (compute_shader.wgsl:65-70)
    pub fn init(&self) -> Result<(), RtError> {
        check(unsafe { ffi::rt_init(self.raw) })
    }

    /// one `update` dispatch (compute_shader.wgsl:333-364) with this frame's 176-byte uniform
    pub fn update(&self, camera: &SceneCamera) -> Result<(), RtError> {
        check(unsafe { ffi::rt_update(self.raw, camera) })
    }

    /// `cameras.len()` consecutive dispatches (the progressive loop; random_seed re-rolled per frame by the caller)
    pub fn update_n(&self, cameras: &[SceneCamera]) -> Result<(), RtError> {
        check(unsafe { ffi::rt_update_n(self.raw, cameras.as_ptr(), cameras.len() as u32) })
    }

    pub fn set_precision(&self, p: Precision) -> Result<(), RtError> {
        let v = match p {
            Precision::Exact => ffi::RT_PRECISION_EXACT,
            Precision::FastShading => ffi::RT_PRECISION_FAST_SHADING,
            Precision::FastAll => ffi::RT_PRECISION_FAST_ALL,
        };
        check(unsafe { ffi::rt_set_precision(self.raw, v) })
    }

    fn texels(&self) -> usize {
        self.width as usize * self.row_count as usize
    }

    /// (mean.rgb, n) per texel of the window, row-major, origin top-left: the reference's Rgba32Float texel
    pub fn read_rgba32f(&self, out: &mut [f32]) -> Result<(), RtError> {
        assert_eq!(out.len(), self.texels() * 4);
        check(unsafe { ffi::rt_read_accum(self.raw, out.as_mut_ptr()) })
    }

    pub fn write_rgba32f(&self, data: &[f32]) -> Result<(), RtError> {
        assert_eq!(data.len(), self.texels() * 4);
        check(unsafe { ffi::rt_write_accum(self.raw, data.as_ptr()) })
    }

    /// linear RGBA8 of the running mean, q(x) = floor(clamp(x, 0, 1) * 255 + 0.5), alpha 255
    pub fn read_rgba8(&self, out: &mut [u8]) -> Result<(), RtError> {
        assert_eq!(out.len(), self.texels() * 4);
        check(unsafe { ffi::rt_read_rgba8(self.raw, out.as_mut_ptr()) })
    }

    /// Frame pipelining: the read-back of this frame overlaps the next update.  `out` must stay alive and untouched
    /// until `read_wait` returns (hence `unsafe`); pinned memory makes the copy asynchronous.
    pub unsafe fn read_rgba8_async(&self, out: *mut u8) -> Result<(), RtError> {
        check(ffi::rt_read_rgba8_async(self.raw, out))
    }

    pub fn read_wait(&self) -> Result<(), RtError> {
        check(unsafe { ffi::rt_read_wait(self.raw) })
    }

    pub fn set_rgba8_output(&self, enable: bool) -> Result<(), RtError> {
        check(unsafe { ffi::rt_set_rgba8_output(self.raw, enable as i32) })
    }

    pub fn synchronize(&self) -> Result<(), RtError> {
        check(unsafe { ffi::rt_synchronize(self.raw) })
    }

    /// sphere_list_hit traversals since the last reset: the "rays" of Mrays/s
    pub fn rays_traced(&self) -> Result<u64, RtError> {
        let mut v = 0u64;
        check(unsafe { ffi::rt_rays_traced(self.raw, &mut v) })?;
        Ok(v)
    }

    pub fn raw(&self) -> *mut ffi::RtRenderer {
        self.raw
    }
}

impl Drop for Renderer {
    fn drop(&mut self) {
        unsafe {
            ffi::rt_destroy(self.raw);
        }
    }
}

/// The headless entry the port target asks for: render(scene, camera, w, h) -> rgba32f (w * h * 4 floats).
pub fn render(spheres: &[GpuSphere], count: u32, camera: &SceneCamera, width: u32, height: u32) -> Result<Vec<f32>, RtError> {
    let mut out = vec![0f32; width as usize * height as usize * 4];
    check(unsafe { ffi::rt_render(0, spheres.as_ptr(), count, camera, width, height, 1, out.as_mut_ptr()) })?;
    Ok(out)
}
