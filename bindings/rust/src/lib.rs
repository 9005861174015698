//! Raw bindings to `include/sfp.h` (libsfp_b200.so): the C ABI that takes the place of the
//! reference engine's CPU physics step.  One `#[repr(C)]` struct per POD of the header, one
//! `extern "C"` declaration per product entry point.  A safe `Physics` wrapper follows below.
//!
//! NOT COMPILED in this repository's environment (no Rust toolchain in the image) and the
//! reference-side names it mirrors (`Physics`, `tick`) are BASELINE.json's vocabulary, not
//! checked against the reference tree (not mounted).  `tests/test_rust_binding.py` checks that
//! every entry point and every struct of the header appears here with the same field list.
#![allow(non_camel_case_types)]

use std::os::raw::c_char;

pub const SFP_OK: i32 = 0;
pub const SFP_E_INVALID: i32 = -1;
pub const SFP_E_CUDA: i32 = -2;
pub const SFP_E_CAPACITY: i32 = -3;
pub const SFP_E_NOMEM: i32 = -5;

pub const SFP_SHAPE_CIRCLE: u32 = 0;
pub const SFP_SHAPE_CAPSULE: u32 = 1;
pub const SFP_SHAPE_RECT: u32 = 2;
pub const SFP_SHAPE_POLYGON: u32 = 3; // | table index << 8
pub const SFP_BODY_STATIC: u32 = 1;
pub const SFP_JOINT_DISTANCE: u32 = 0;
pub const SFP_JOINT_ROPE: u32 = 1;
pub const SFP_FLAG_BROADPHASE_EVERY_SUBSTEP: u32 = 1;
pub const SFP_FLAG_MULTI_KERNEL: u32 = 2;
pub const SFP_FLAG_FLOW_SOLVE: u32 = 4;
pub const SFP_FLAG_TILES: u32 = 8;
pub const SFP_FLAG_BATCH_WHOLE_WORLDS: u32 = 16;

#[repr(C)]
#[derive(Clone, Copy)]
pub struct SfpParams {
    pub gravity_x: f32,
    pub gravity_y: f32,
    pub cell_size: f32,
    pub aabb_margin: f32,
    pub vel_expand: f32,
    pub contact_margin: f32,
    pub solver_order: u32,
    pub flags: u32,
    pub max_bodies: u32,
    pub max_joints: u32,
    pub max_entries: u32,
    pub max_pairs: u32,
    pub device: i32,
    pub reserved: [u32; 3],
} // 64 bytes

#[repr(C)]
#[derive(Clone, Copy, Default)]
pub struct SfpBodyDesc {
    pub x: f32,
    pub y: f32,
    pub angle: f32,
    pub vx: f32,
    pub vy: f32,
    pub omega: f32,
    pub hx: f32,
    pub hy: f32,
    pub radius: f32,
    pub shape: u32,
    pub density: f32,
    pub mu_s: f32,
    pub mu_d: f32,
    pub restitution: f32,
    pub category: u16,
    pub mask: u16,
    pub flags: u32,
} // 64 bytes

#[repr(C)]
#[derive(Clone, Copy, Default)]
pub struct SfpJointDesc {
    pub body_a: u32,
    pub body_b: u32,
    pub ax: f32,
    pub ay: f32,
    pub bx: f32,
    pub by: f32,
    pub rest: f32,
    pub compliance: f32,
    pub kind: u32,
    pub reserved: [u32; 3],
} // 48 bytes

#[repr(C)]
#[derive(Clone, Copy, Default)]
pub struct SfpBodyState {
    pub x: f32,
    pub y: f32,
    pub c: f32,
    pub s: f32,
    pub vx: f32,
    pub vy: f32,
    pub omega: f32,
    pub angle: f32,
} // 32 bytes

#[repr(C)]
#[derive(Clone, Copy)]
pub struct SfpPolygon {
    pub count: u32,
    pub reserved: u32,
    pub x: [f32; 8],
    pub y: [f32; 8],
} // 72 bytes

#[repr(C)]
#[derive(Clone, Copy, Default)]
pub struct SfpPair {
    pub a: u32,
    pub b: u32,
}

#[repr(C)]
#[derive(Clone, Copy, Default)]
pub struct SfpContact {
    pub a: u32,
    pub b: u32,
    pub count: u32,
    pub nx: f32,
    pub ny: f32,
    pub depth: [f32; 2],
    pub px: [f32; 2],
    pub py: [f32; 2],
    pub lambda_n: [f32; 2],
    pub lambda_t: [f32; 2],
} // 60 bytes

#[repr(C)]
#[derive(Clone, Copy, Default)]
pub struct SfpContactEvent {
    pub a: u32,
    pub b: u32,
    pub nx: f32,
    pub ny: f32,
    pub px: f32,
    pub py: f32,
    pub impulse: f32,
    pub points: u32,
} // 32 bytes

#[repr(C)]
#[derive(Clone, Copy, Default)]
pub struct SfpRay {
    pub ox: f32,
    pub oy: f32,
    pub dx: f32,
    pub dy: f32,
    pub max_t: f32,
    pub mask: u32,
    pub reserved: [u32; 2],
} // 32 bytes

#[repr(C)]
#[derive(Clone, Copy, Default)]
pub struct SfpRayHit {
    pub body: u32,
    pub t: f32,
    pub nx: f32,
    pub ny: f32,
} // 16 bytes

#[repr(C)]
#[derive(Clone, Copy, Default)]
pub struct SfpStats {
    pub frames: u64,
    pub body_substeps: u64,
    pub n_bodies: u32,
    pub n_dynamic: u32,
    pub n_joints: u32,
    pub n_entries: u32,
    pub n_pairs: u32,
    pub n_contacts: u32,
    pub n_manifolds: u32,
    pub n_colours: u32,
    pub colour_rounds: u32,
    pub overflow: u32,
    pub max_penetration: f32,
    pub kinetic_energy: f32,
    pub potential_energy: f32,
    pub launches: u32,
    pub reserved: [u32; 2],
} // 80 bytes

#[repr(C)]
pub struct SfpWorld {
    _private: [u8; 0],
}
#[repr(C)]
pub struct SfpBatch {
    _private: [u8; 0],
}
#[repr(C)]
pub struct SfpBatchMulti {
    _private: [u8; 0],
}

extern "C" {
    // ---- single world ----
    pub fn sfp_default_params(out: *mut SfpParams) -> i32;
    pub fn sfp_create(params: *const SfpParams, out: *mut *mut SfpWorld) -> i32;
    pub fn sfp_destroy(w: *mut SfpWorld) -> i32;
    pub fn sfp_upload_bodies(w: *mut SfpWorld, bodies: *const SfpBodyDesc, n: u32) -> i32;
    pub fn sfp_set_polygons(w: *mut SfpWorld, polys: *const SfpPolygon, n: u32) -> i32;
    pub fn sfp_set_joints(w: *mut SfpWorld, joints: *const SfpJointDesc, n: u32) -> i32;
    pub fn sfp_step(w: *mut SfpWorld, dt: f32, substeps: u32) -> i32;
    pub fn sfp_sync(w: *mut SfpWorld) -> i32;
    pub fn sfp_download_bodies(w: *mut SfpWorld, out: *mut SfpBodyState, n: u32) -> i32;
    pub fn sfp_get_pairs(w: *mut SfpWorld, out: *mut SfpPair, cap: u32, n_out: *mut u32) -> i32;
    pub fn sfp_get_contacts(w: *mut SfpWorld, out: *mut SfpContact, cap: u32, n_out: *mut u32) -> i32;
    pub fn sfp_get_colours(w: *mut SfpWorld, pair_colours: *mut u32, cap: u32,
                           joint_colours: *mut u32, jcap: u32) -> i32;
    pub fn sfp_get_stats(w: *mut SfpWorld, out: *mut SfpStats) -> i32;
    pub fn sfp_last_error(w: *mut SfpWorld) -> *const c_char;
    pub fn sfp_get_contact_events(w: *mut SfpWorld, out: *mut SfpContactEvent, cap: u32,
                                  n_out: *mut u32) -> i32;
    pub fn sfp_ray_cast(w: *mut SfpWorld, rays: *const SfpRay, m: u32, hits: *mut SfpRayHit) -> i32;
    pub fn sfp_query_aabb(w: *mut SfpWorld, lo_x: f32, lo_y: f32, hi_x: f32, hi_y: f32, mask: u32,
                          out: *mut u32, cap: u32, n_out: *mut u32) -> i32;
    pub fn sfp_query_shape(w: *mut SfpWorld, probe: *const SfpBodyDesc, mask: u32, out: *mut u32,
                           cap: u32, n_out: *mut u32) -> i32;
    pub fn sfp_update_bodies(w: *mut SfpWorld, idx: *const u32, states: *const SfpBodyState,
                             m: u32) -> i32;
    pub fn sfp_step_host(w: *mut SfpWorld, input: *const SfpBodyState, out: *mut SfpBodyState,
                         n: u32, dt: f32, substeps: u32) -> i32;
    pub fn sfp_alloc_host_states(n: u32, out: *mut *mut SfpBodyState) -> i32;
    pub fn sfp_free_host_states(p: *mut SfpBodyState) -> i32;
    pub fn sfp_last_step_ms(w: *mut SfpWorld, ms: *mut f32) -> i32;
    pub fn sfp_last_step_breakdown_ms(w: *mut SfpWorld, out: *mut f32) -> i32; // float[4]
    pub fn sfp_last_step_stage_ns(w: *mut SfpWorld, out: *mut u64) -> i32; // uint64_t[6]
    pub fn sfp_backend() -> *const c_char;
    // ---- batched independent worlds ----
    pub fn sfp_batch_create(params: *const SfpParams, n_worlds: u32, bodies_per_world: u32,
                            out: *mut *mut SfpBatch) -> i32;
    pub fn sfp_batch_destroy(b: *mut SfpBatch) -> i32;
    pub fn sfp_batch_upload(b: *mut SfpBatch, bodies: *const SfpBodyDesc) -> i32;
    pub fn sfp_batch_set_polygons(b: *mut SfpBatch, polys: *const SfpPolygon, n: u32) -> i32;
    pub fn sfp_batch_set_joints(b: *mut SfpBatch, joints: *const SfpJointDesc, joints_per_world: u32,
                                counts: *const u32) -> i32;
    pub fn sfp_batch_step(b: *mut SfpBatch, dt: f32, substeps: u32) -> i32;
    pub fn sfp_batch_sync(b: *mut SfpBatch) -> i32;
    pub fn sfp_batch_download(b: *mut SfpBatch, out: *mut SfpBodyState) -> i32;
    pub fn sfp_batch_step_host(b: *mut SfpBatch, input: *const SfpBodyState,
                               out: *mut SfpBodyState, dt: f32, substeps: u32) -> i32;
    pub fn sfp_batch_stats(b: *mut SfpBatch, out: *mut SfpStats) -> i32;
    pub fn sfp_batch_last_error(b: *mut SfpBatch) -> *const c_char;
    pub fn sfp_batch_last_step_ms(b: *mut SfpBatch, ms: *mut f32) -> i32;
    // ---- batched worlds over several GPUs from one process ----
    pub fn sfp_batch_create_multi(params: *const SfpParams, n_worlds: u32, bodies_per_world: u32,
                                  device_ids: *const i32, n_dev: u32, out: *mut *mut SfpBatchMulti) -> i32;
    pub fn sfp_batch_multi_destroy(m: *mut SfpBatchMulti) -> i32;
    pub fn sfp_batch_multi_partition(m: *mut SfpBatchMulti, dev_index: u32, device: *mut i32,
                                     first_world: *mut u32, n_worlds: *mut u32) -> i32;
    pub fn sfp_batch_multi_upload(m: *mut SfpBatchMulti, bodies: *const SfpBodyDesc) -> i32;
    pub fn sfp_batch_multi_set_polygons(m: *mut SfpBatchMulti, polys: *const SfpPolygon, n: u32) -> i32;
    pub fn sfp_batch_multi_set_joints(m: *mut SfpBatchMulti, joints: *const SfpJointDesc,
                                      joints_per_world: u32, counts: *const u32) -> i32;
    pub fn sfp_batch_multi_step(m: *mut SfpBatchMulti, dt: f32, substeps: u32) -> i32;
    pub fn sfp_batch_multi_sync(m: *mut SfpBatchMulti) -> i32;
    pub fn sfp_batch_multi_download(m: *mut SfpBatchMulti, out: *mut SfpBodyState) -> i32;
    pub fn sfp_batch_multi_step_host(m: *mut SfpBatchMulti, input: *const SfpBodyState,
                                     out: *mut SfpBodyState, dt: f32, substeps: u32) -> i32;
    pub fn sfp_batch_multi_stats(m: *mut SfpBatchMulti, out: *mut SfpStats) -> i32;
    pub fn sfp_batch_multi_last_step_ms(m: *mut SfpBatchMulti, ms: *mut f32) -> i32;
    pub fn sfp_batch_multi_last_error(m: *mut SfpBatchMulti) -> *const c_char;
    // ---- one small world without any allocation ----
    pub fn sfp_step_small_noalloc(params: *const SfpParams, bodies: *const SfpBodyDesc, n: u32,
                                  dt: f32, substeps: u32, frames: u32, out: *mut SfpBodyState,
                                  err: *mut c_char, err_cap: u32) -> i32;
}

// ---------------------------------------------------------------------------------------------
// The safe wrapper that takes the place of the CPU step behind the engine's `Physics` type.
// ---------------------------------------------------------------------------------------------
#[derive(Debug)]
pub struct PhysicsError {
    pub code: i32,
    pub detail: String,
}

fn check(code: i32, w: *mut SfpWorld) -> Result<(), PhysicsError> {
    if code == SFP_OK {
        return Ok(());
    }
    let detail = unsafe {
        let p = sfp_last_error(w);
        if p.is_null() { String::new() } else { std::ffi::CStr::from_ptr(p).to_string_lossy().into_owned() }
    };
    Err(PhysicsError { code, detail })
}

/// `&mut self` everywhere: a world is not thread-safe, exactly like the CPU `Physics` it replaces.
pub struct Physics {
    world: *mut SfpWorld,
    substeps: u32,
    descs: Vec<SfpBodyDesc>,
    dirty: bool,
}

impl Physics {
    pub fn new(gravity: [f32; 2], substeps: u32) -> Result<Self, PhysicsError> {
        let mut p = std::mem::MaybeUninit::<SfpParams>::uninit();
        unsafe { sfp_default_params(p.as_mut_ptr()) };
        let mut p = unsafe { p.assume_init() };
        p.gravity_x = gravity[0];
        p.gravity_y = gravity[1];
        let mut world = std::ptr::null_mut();
        check(unsafe { sfp_create(&p, &mut world) }, world)?;
        Ok(Self { world, substeps, descs: Vec::new(), dirty: true })
    }

    /// scene-graph walk: one desc per (Body, Collider) pair, in a stable order
    pub fn set_bodies(&mut self, descs: impl Iterator<Item = SfpBodyDesc>) {
        self.descs.clear();
        self.descs.extend(descs);
        self.dirty = true;
    }

    pub fn set_joints(&mut self, joints: &[SfpJointDesc]) -> Result<(), PhysicsError> {
        check(unsafe { sfp_set_joints(self.world, joints.as_ptr(), joints.len() as u32) }, self.world)
    }

    /// the per-frame step (asynchronous until `write_back`)
    pub fn tick(&mut self, dt: f32) -> Result<(), PhysicsError> {
        if self.dirty {
            check(unsafe { sfp_upload_bodies(self.world, self.descs.as_ptr(), self.descs.len() as u32) },
                  self.world)?;
            self.dirty = false;
        }
        check(unsafe { sfp_step(self.world, dt, self.substeps) }, self.world)
    }

    /// pose write-back into the scene graph (synchronises)
    pub fn write_back(&mut self, out: &mut [SfpBodyState]) -> Result<(), PhysicsError> {
        check(unsafe { sfp_download_bodies(self.world, out.as_mut_ptr(), out.len() as u32) }, self.world)
    }

    /// one call per frame for engines that keep body state on the host: states in, states out
    pub fn tick_host(&mut self, input: &[SfpBodyState], out: &mut [SfpBodyState], dt: f32)
                     -> Result<(), PhysicsError> {
        assert_eq!(input.len(), out.len());
        check(unsafe { sfp_step_host(self.world, input.as_ptr(), out.as_mut_ptr(), out.len() as u32,
                                     dt, self.substeps) }, self.world)
    }

    pub fn contact_events(&mut self, out: &mut Vec<SfpContactEvent>) -> Result<(), PhysicsError> {
        let mut n = 0u32;
        check(unsafe { sfp_get_contact_events(self.world, std::ptr::null_mut(), 0, &mut n) }, self.world)?;
        out.resize(n as usize, SfpContactEvent::default());
        check(unsafe { sfp_get_contact_events(self.world, out.as_mut_ptr(), n, &mut n) }, self.world)
    }

    pub fn ray_cast(&mut self, rays: &[SfpRay], hits: &mut [SfpRayHit]) -> Result<(), PhysicsError> {
        assert_eq!(rays.len(), hits.len());
        check(unsafe { sfp_ray_cast(self.world, rays.as_ptr(), rays.len() as u32, hits.as_mut_ptr()) },
              self.world)
    }

    /// bodies whose collider overlaps `probe` (shape, x, y, angle, hx, hy, radius of a desc; the
    /// narrowphase with zero margin decides); a circle of radius 0 is a point query
    pub fn query_shape(&mut self, probe: &SfpBodyDesc, mask: u32, out: &mut Vec<u32>) -> Result<(), PhysicsError> {
        let mut n = 0u32;
        check(unsafe { sfp_query_shape(self.world, probe, mask, std::ptr::null_mut(), 0, &mut n) }, self.world)?;
        out.resize(n as usize, 0);
        check(unsafe { sfp_query_shape(self.world, probe, mask, out.as_mut_ptr(), n, &mut n) }, self.world)
    }
}

impl Drop for Physics {
    fn drop(&mut self) {
        unsafe { sfp_destroy(self.world) };
    }
}
