//! Rust binding of include/relayrl_b200.h for relayrl_framework (NOT compiled in this repository:
//! the build image has no Rust toolchain).  Drop into relayrl_framework/src/native/b200.rs and add
//! `println!("cargo:rustc-link-lib=dylib=relayrl_b200");` to build.rs.  See INTEGRATION.md section 2 for
//! the call sites (training_zmq.rs:994-1029, training_grpc.rs:585-599, agent_zmq.rs:481).
use std::ffi::{c_char, c_void, CStr};

#[repr(C)]
pub struct RrlConfig {
    pub obs_dim: i32, pub act_dim: i32, pub hidden: i32, pub with_baseline: i32, pub train_vf_iters: i32,
    pub logp_mode: i32, pub device: i32, pub rank: i32, pub nranks: i32, pub flags: i32,
    pub capacity: i64, pub gamma: f64, pub lam: f64, pub pi_lr: f64, pub vf_lr: f64,
    pub adam_beta1: f64, pub adam_beta2: f64, pub adam_eps: f64, pub nccl_id: [u8; 128],
}

#[repr(C)]
pub struct RrlBatchView {
    pub obs: *const c_void, pub act: *const f32, pub mask: *const c_void, pub rew: *const f32,
    pub val: *const f32, pub logp: *const f32, pub done: *const u8, pub count: i64,
    pub obs_dtype: i32, pub mask_dtype: i32,
}

#[repr(C)]
#[derive(Default)]
pub struct RrlEpochStats {
    pub loss_pi: f64, pub delta_loss_pi: f64, pub kl: f64, pub entropy: f64,
    pub loss_v: f64, pub delta_loss_v: f64, pub loss_v_old: f64, pub adv_mean: f64, pub adv_std: f64,
    pub n_transitions: i64, pub n_episodes: i64,
    pub ms_scan: f32, pub ms_policy: f32, pub ms_value: f32, pub ms_total: f32,
    pub kernel_launches: i32, pub reserved: i32,
}

extern "C" {
    pub fn rrl_config_defaults(cfg: *mut RrlConfig);
    pub fn rrl_learner_create(cfg: *const RrlConfig, out: *mut *mut c_void) -> i32;
    pub fn rrl_learner_destroy(h: *mut c_void) -> i32;
    pub fn rrl_learner_stream(h: *const c_void, stream: *mut *mut c_void) -> i32;
    pub fn rrl_ingest_grpc_trajectory(h: *mut c_void, msg: *const u8, len: usize, n_actions: *mut i64) -> i32;
    pub fn rrl_set_params(h: *mut c_void, which: i32, flat: *const f32, n: i64) -> i32;
    pub fn rrl_get_params(h: *mut c_void, which: i32, flat: *mut f32, n: i64) -> i32;
    pub fn rrl_ingest_packed(h: *mut c_void, batch: *const RrlBatchView) -> i32;
    pub fn rrl_ingest_wire_action(h: *mut c_void, obs: *const u8, obs_len: usize, act: *const u8, act_len: usize,
                                  mask: *const u8, mask_len: usize, rew: f32, v: *const u8, v_len: usize,
                                  logp: *const u8, logp_len: usize, done: u8) -> i32;
    pub fn rrl_train_epoch(h: *mut c_void, stats: *mut RrlEpochStats) -> i32;
    pub fn rrl_last_error(h: *const c_void) -> *const c_char;
    // the ZMQ transport's message body: serde_pickle of Vec<RelayRLAction> (training_zmq.rs:994-1001), as received
    pub fn rrl_ingest_pickle(h: *mut c_void, msg: *const u8, len: usize, n_actions: *mut i64) -> i32;
    // weights, Adam state and the buffer's stale adv/ret tail (the reference keeps none of it across a restart)
    pub fn rrl_checkpoint_save(h: *mut c_void, path: *const c_char, user_tag: u64) -> i32;
    pub fn rrl_checkpoint_load(h: *mut c_void, path: *const c_char, user_tag: *mut u64) -> i32;
    // batched agent: request_for_action / flag_last_action for n_envs environments per call (agent_zmq.rs:458-610)
    pub fn rrl_actor_create(obs_dim: i32, act_dim: i32, hidden: i32, with_baseline: i32, device: i32, max_batch: i64,
                            out: *mut *mut c_void) -> i32;
    pub fn rrl_actor_set_params(a: *mut c_void, which: i32, flat: *const f32, n: i64) -> i32;
    pub fn rrl_actor_step(a: *mut c_void, obs: *const f32, mask: *const f32, batch: i64, seed: u64, offset: u64,
                          act: *mut i64, logp_a: *mut f32, v: *mut f32, probs: *mut f32, logits: *mut f32) -> i32;
    pub fn rrl_rollout_create(actor: *mut c_void, learner: *mut c_void, n_envs: i64, max_len: i64, out: *mut *mut c_void) -> i32;
    pub fn rrl_rollout_step(r: *mut c_void, obs: *const f32, mask: *const f32, reward: *const f32, seed: u64, offset: u64,
                            act: *mut i64, logp_a: *mut f32, v: *mut f32) -> i32;
    pub fn rrl_rollout_flag_last(r: *mut c_void, env_ids: *const i64, rewards: *const f32, count: i64) -> i32;
    pub fn rrl_rollout_destroy(r: *mut c_void) -> i32;
}

/// Safe wrapper: one learner = one GPU; `&mut self` serialises calls like the worker's asyncio.Lock.
pub struct B200Learner(*mut c_void);
unsafe impl Send for B200Learner {}

impl B200Learner {
    pub fn new(mut cfg: RrlConfig) -> Result<Self, String> {
        let mut h = std::ptr::null_mut();
        let st = unsafe { rrl_learner_create(&mut cfg, &mut h) };
        if st != 0 { return Err(last_error(std::ptr::null())); }
        Ok(Self(h))
    }
    /// RelayRLAction with TensorData blobs (src/types/action.rs:428-437) -> device buffers.
    pub fn ingest_action(&mut self, obs: &[u8], act: &[u8], mask: &[u8], rew: f32, v: &[u8], logp: &[u8], done: bool)
        -> Result<(), String> {
        let st = unsafe { rrl_ingest_wire_action(self.0, obs.as_ptr(), obs.len(), act.as_ptr(), act.len(),
            mask.as_ptr(), mask.len(), rew, v.as_ptr(), v.len(), logp.as_ptr(), logp.len(), done as u8) };
        if st != 0 { Err(last_error(self.0)) } else { Ok(()) }
    }
    /// The ZMQ server's loop body: one pickled trajectory as it came off the socket.
    pub fn ingest_pickle(&mut self, msg: &[u8]) -> Result<i64, String> {
        let mut n = 0i64;
        let st = unsafe { rrl_ingest_pickle(self.0, msg.as_ptr(), msg.len(), &mut n) };
        if st != 0 { Err(last_error(self.0)) } else { Ok(n) }
    }
    pub fn train_epoch(&mut self) -> Result<RrlEpochStats, String> {
        let mut s = RrlEpochStats::default();
        let st = unsafe { rrl_train_epoch(self.0, &mut s) };
        if st != 0 { Err(last_error(self.0)) } else { Ok(s) }
    }
}
impl Drop for B200Learner { fn drop(&mut self) { unsafe { rrl_learner_destroy(self.0); } } }

fn last_error(h: *const c_void) -> String {
    unsafe { CStr::from_ptr(rrl_last_error(h)).to_string_lossy().into_owned() }
}
