//! moose/src/host/b200.rs -- the shim a Moose maintainer adds to run the `[runtime]` HostPlacement kernels on
//! libmoose_b200.so (feature `b200`).  SOURCE ONLY: this image has no Rust toolchain (no cargo / rustc), so the file
//! has never been compiled; the same entry points are exercised through Python ctypes by the parity tests
//! (moose_b200/__init__.py, tests/).  The `extern "C"` block it relies on is generated from the header
//! (integration/rust/b200_ffi.rs, `python tools/gen_rust_ffi.py`).
//!
//! What it replaces, file:line in plurigrid/moose:
//!   * `HostRingTensor<T>` gets a device-backed sibling, `DeviceRing<T>` (moose/src/host/mod.rs:388), and
//!     `HostBitTensor` a `DeviceBits` (moose/src/host/mod.rs:205); both convert from / to `Value`
//!     (moose/src/computation.rs:390-400) so the `Kernel::{Nullary..Variadic}` closures of
//!     moose/src/kernels/mod.rs:34-106 take them unchanged.
//!   * the bodies of the `[runtime]` kernels of moose/src/host/ops.rs and moose/src/host/prim.rs listed below.
//!   * serde of device tensors for `Send` / `Receive` (moose/src/networking/mod.rs:16-54): bytes leave the device
//!     once, in the layout ndarray serialises (row-major, little endian).
//!
//! Ownership follows the reference: operands by value (cheap clones of an `Arc`), a NEW tensor per result, no
//! aliasing; the placement is re-stamped on the result (`plc.clone()`), cross-host placement stays a relabel.
#![cfg(feature = "b200")]

use std::ffi::CStr;
use std::marker::PhantomData;
use std::num::Wrapping;
use std::sync::Arc;

use ndarray::{ArrayD, IxDyn};
use serde::{Deserialize, Deserializer, Serialize, Serializer};

use crate::computation::{HostPlacement, RendezvousKey, SessionId, SyncKey, Value};
use crate::error::{Error, Result};
use crate::execution::RuntimeSession;
use crate::host::b200_ffi::*;
use crate::host::{HostBitTensor, HostPrfKey, HostRingTensor, HostSeed, HostShape, RawPrfKey, RawSeed, RawShape};

pub const MB_RING64: i32 = 0;
pub const MB_RING128: i32 = 1;
pub const MB_BIT: i32 = 2;
pub const MB_FLOAT64: i32 = 3;

// ---- context -----------------------------------------------------------------------------------------------------------

/// One library context per (session, device): its stream orders the session's kernels, its session id feeds
/// DeriveSeed (moose/src/host/prim.rs:130).  Sessions keep it in their state (`SyncSession` / `AsyncSession` gain a
/// field); the library is thread safe, so the tokio workers of `AsyncSession`
/// (moose/src/execution/asynchronous.rs:453-531) share it.
pub struct B200Context(pub(crate) *mut MbCtx);
unsafe impl Send for B200Context {}
unsafe impl Sync for B200Context {}

impl B200Context {
    pub fn new(device: i32, session_id: &SessionId) -> Result<Self> {
        let mut ctx = std::ptr::null_mut();
        check(std::ptr::null_mut(), unsafe { mb_ctx_create(device, &mut ctx) })?;
        check(ctx, unsafe { mb_ctx_set_session_raw(ctx, session_id.as_bytes().as_ptr()) })?;
        Ok(B200Context(ctx))
    }
}
impl Drop for B200Context {
    fn drop(&mut self) {
        unsafe { mb_ctx_destroy(self.0) };
    }
}

/// Sessions that can run device kernels hand out their context.
pub trait B200Session: RuntimeSession {
    fn b200(&self) -> &B200Context;
}

pub(crate) fn check(ctx: *mut MbCtx, rc: i32) -> Result<()> {
    if rc == 0 {
        return Ok(());
    }
    let msg = unsafe { CStr::from_ptr(mb_last_error(ctx)) }.to_string_lossy().into_owned();
    Err(match rc {
        -1 => Error::InvalidArgument(msg),          // MB_ERR_INVALID_ARGUMENT
        -3 => Error::UnimplementedOperator(msg),    // MB_ERR_UNIMPLEMENTED
        -4 => Error::TypeMismatch { expected: msg, found: crate::computation::Ty::Unknown },
        _ => Error::KernelError(msg),               // MB_ERR_KERNEL, MB_ERR_CUDA, MB_ERR_NO_DEVICE
    })
}

// ---- device tensors ------------------------------------------------------------------------------------------------------

pub(crate) struct DeviceHandle {
    ctx: *mut MbCtx,
    t: *mut MbTensor,
}
unsafe impl Send for DeviceHandle {}
unsafe impl Sync for DeviceHandle {}
impl Drop for DeviceHandle {
    fn drop(&mut self) {
        unsafe { mb_tensor_free(self.ctx, self.t) };   // stream-ordered free
    }
}

/// Width of the ring element behind `T` (u64 / u128), as the library's dtype tag.
pub trait RingWidth: Copy + 'static {
    const DTYPE: i32;
}
impl RingWidth for u64 {
    const DTYPE: i32 = MB_RING64;
}
impl RingWidth for u128 {
    const DTYPE: i32 = MB_RING128;
}

/// Device-resident sibling of `HostRingTensor<T>` (moose/src/host/mod.rs:388): same placement tag, same value
/// semantics (immutable, shared by reference count), contents in the GPU's HBM.
#[derive(Clone)]
pub struct DeviceRing<T>(pub(crate) Arc<DeviceHandle>, pub HostPlacement, PhantomData<T>);

/// Device-resident sibling of `HostBitTensor` (moose/src/host/mod.rs:205): packed Lsb0 like `BitVec<u8, Lsb0>`.
#[derive(Clone)]
pub struct DeviceBits(pub(crate) Arc<DeviceHandle>, pub HostPlacement);

impl<T: RingWidth> DeviceRing<T> {
    pub(crate) fn from_raw(ctx: *mut MbCtx, t: *mut MbTensor, plc: HostPlacement) -> Self {
        DeviceRing(Arc::new(DeviceHandle { ctx, t }), plc, PhantomData)
    }
    pub(crate) fn raw(&self) -> *const MbTensor {
        self.0.t
    }
    pub fn shape(&self) -> Vec<usize> {
        let rank = unsafe { mb_tensor_rank(self.0.t) } as usize;
        let dims = unsafe { std::slice::from_raw_parts(mb_tensor_shape(self.0.t), rank) };
        dims.iter().map(|d| *d as usize).collect()
    }
    /// Upload of a host tensor (edge of the graph: Input / Constant / Load results).
    pub fn from_host<S: B200Session>(sess: &S, x: &HostRingTensor<T>) -> Result<Self>
    where
        Wrapping<T>: Clone,
    {
        let ctx = sess.b200().0;
        let shape: Vec<i64> = x.0.shape().iter().map(|d| *d as i64).collect();
        let std_layout = x.0.as_standard_layout();           // ndarray views may be strided (transpose)
        let mut out = std::ptr::null_mut();
        check(ctx, unsafe {
            mb_tensor_from_host(ctx, T::DTYPE, shape.len() as i32, shape.as_ptr(), std_layout.as_ptr() as *const _, &mut out)
        })?;
        Ok(Self::from_raw(ctx, out, x.1.clone()))
    }
    /// Download (Output / Save / Send to a peer that is not on this box).
    pub fn to_host(&self) -> Result<HostRingTensor<T>>
    where
        T: Default,
    {
        let shape = self.shape();
        let n: usize = shape.iter().product();
        let mut buf: Vec<Wrapping<T>> = (0..n).map(|_| Wrapping(T::default())).collect();
        check(self.0.ctx, unsafe {
            mb_tensor_to_host(self.0.ctx, self.0.t, buf.as_mut_ptr() as *mut _, n * std::mem::size_of::<T>())
        })?;
        let arr = ArrayD::from_shape_vec(IxDyn(&shape), buf).map_err(|e| Error::KernelError(e.to_string()))?;
        Ok(HostRingTensor(arr.into_shared(), self.1.clone()))
    }
}

impl DeviceBits {
    pub(crate) fn from_raw(ctx: *mut MbCtx, t: *mut MbTensor, plc: HostPlacement) -> Self {
        DeviceBits(Arc::new(DeviceHandle { ctx, t }), plc)
    }
    pub(crate) fn raw(&self) -> *const MbTensor {
        self.0.t
    }
}

// `Value` gains `DeviceRing64Tensor(Box<DeviceRing<u64>>)`, `DeviceRing128Tensor(..)`, `DeviceBitTensor(..)` variants
// through the `values!` macro of moose/src/computation.rs (one line each); TryFrom / Into follow the pattern the macro
// generates for the host variants, so `Kernel::Binary` closures (moose/src/kernels/mod.rs:60-75) bind to them as is:
impl<T: RingWidth> TryFrom<Value> for DeviceRing<T>
where
    Value: DeviceRingValue<T>,
{
    type Error = Error;
    fn try_from(v: Value) -> Result<Self> {
        <Value as DeviceRingValue<T>>::into_device(v)
    }
}
/// Implemented for u64 / u128 next to the generated `Value` variants.
pub trait DeviceRingValue<T> {
    fn into_device(self) -> Result<DeviceRing<T>>;
    fn from_device(x: DeviceRing<T>) -> Self;
}

// ---- serde: Send / Receive and storage (moose/src/networking/mod.rs:16-54) ---------------------------------------------------
// A device tensor crosses a process boundary as the host tensor it stands for: bincode / msgpack see exactly the
// bytes they see today (ndarray's shape + row-major elements), so a peer without a GPU -- or the reference's own
// gRPC / TCP networking (moose/src/networking/{grpc,tcpstream}.rs) -- interoperates.  Peers that are both on B200s of one
// box skip this path: the party runtime moves reshares over NVLink (csrc/peer_transport.cu).
impl<T: RingWidth + Default + Serialize> Serialize for DeviceRing<T>
where
    Wrapping<T>: Clone,
{
    fn serialize<Ser: Serializer>(&self, s: Ser) -> std::result::Result<Ser::Ok, Ser::Error> {
        self.to_host().map_err(serde::ser::Error::custom)?.serialize(s)
    }
}
/// Deserialisation yields the host tensor; the first device kernel that takes it uploads it (`ensure_device`), because
/// a deserialiser has no session, hence no context / device, at hand.
pub enum RingOperand<T> {
    Host(HostRingTensor<T>),
    Device(DeviceRing<T>),
}
impl<'de, T: Deserialize<'de>> Deserialize<'de> for RingOperand<T>
where
    HostRingTensor<T>: Deserialize<'de>,
{
    fn deserialize<D: Deserializer<'de>>(d: D) -> std::result::Result<Self, D::Error> {
        Ok(RingOperand::Host(HostRingTensor::<T>::deserialize(d)?))
    }
}
impl<T: RingWidth> RingOperand<T>
where
    Wrapping<T>: Clone,
{
    pub fn ensure_device<S: B200Session>(self, sess: &S) -> Result<DeviceRing<T>> {
        match self {
            RingOperand::Device(x) => Ok(x),
            RingOperand::Host(x) => DeviceRing::from_host(sess, &x),
        }
    }
}

/// `SyncNetworking::send` for a value that lives on the device (moose/src/networking/mod.rs:16-24): download once, hand
/// the host value to the existing transport.
pub fn send_device_ring<N: crate::networking::SyncNetworking, T: RingWidth + Default>(
    net: &N,
    x: &DeviceRing<T>,
    receiver: &crate::execution::Identity,
    rendezvous_key: &RendezvousKey,
    session_id: &SessionId,
) -> Result<()>
where
    Value: From<HostRingTensor<T>>,
{
    net.send(&Value::from(x.to_host()?), receiver, rendezvous_key, session_id)
}

// ---- kernel bodies ---------------------------------------------------------------------------------------------------------

macro_rules! ring_binary {
    ($op:ident, $ffi:ident, $cite:literal) => {
        impl crate::computation::$op {
            #[doc = $cite]
            pub(crate) fn ring_kernel<S: B200Session, T: RingWidth>(
                sess: &S,
                plc: &HostPlacement,
                x: DeviceRing<T>,
                y: DeviceRing<T>,
            ) -> Result<DeviceRing<T>> {
                let ctx = sess.b200().0;
                let mut out = std::ptr::null_mut();
                check(ctx, unsafe { $ffi(ctx, x.raw(), y.raw(), &mut out) })?;   // ndarray co-broadcasting included
                Ok(DeviceRing::from_raw(ctx, out, plc.clone()))
            }
        }
    };
}
ring_binary!(AddOp, mb_add, "replaces moose/src/host/ops.rs:1709-1722");
ring_binary!(SubOp, mb_sub, "replaces moose/src/host/ops.rs:1724-1737");
ring_binary!(MulOp, mb_mul, "replaces moose/src/host/ops.rs:299-310");
ring_binary!(DotOp, mb_dot, "replaces moose/src/host/ops.rs:1754-1851 (ranks (1,1), (1,2), (2,1), (2,2); other ranks are KernelError like the reference)");
ring_binary!(DivOp, mb_div, "replaces moose/src/host/ops.rs:312-324");

impl crate::computation::NegOp {
    /// replaces moose/src/host/ops.rs:1739-1752
    pub(crate) fn ring_kernel<S: B200Session, T: RingWidth>(sess: &S, plc: &HostPlacement, x: DeviceRing<T>) -> Result<DeviceRing<T>> {
        let ctx = sess.b200().0;
        let mut out = std::ptr::null_mut();
        check(ctx, unsafe { mb_neg(ctx, x.raw(), &mut out) })?;
        Ok(DeviceRing::from_raw(ctx, out, plc.clone()))
    }
}

macro_rules! ring_shift {
    ($op:ident, $ffi:ident, $cite:literal) => {
        impl crate::computation::$op {
            #[doc = $cite]
            pub(crate) fn ring_kernel<S: B200Session, T: RingWidth>(
                sess: &S,
                plc: &HostPlacement,
                amount: usize,
                x: DeviceRing<T>,
            ) -> Result<DeviceRing<T>> {
                let ctx = sess.b200().0;
                let mut out = std::ptr::null_mut();
                // the library masks the amount mod the bit width, like `Wrapping<T> << usize`
                check(ctx, unsafe { $ffi(ctx, amount as u64, x.raw(), &mut out) })?;
                Ok(DeviceRing::from_raw(ctx, out, plc.clone()))
            }
        }
    };
}
ring_shift!(ShlOp, mb_shl, "replaces moose/src/host/ops.rs:1853-1866");
ring_shift!(ShrOp, mb_shr, "replaces moose/src/host/ops.rs:1868-1881");

fn shape_i64(shape: &HostShape) -> Vec<i64> {
    shape.0 .0.iter().map(|d| *d as i64).collect()
}

impl crate::computation::SampleSeededOp {
    /// replaces moose/src/host/ops.rs:1959-1988 (ring64) and :1990-2017 (ring128): the AES-128-CTR stream of
    /// `AesRng::from_seed`, element order as the reference; `max_value` other than None / Some(1) is
    /// UnimplementedOperator like the reference.
    pub(crate) fn ring_kernel<S: B200Session, T: RingWidth>(
        sess: &S,
        plc: &HostPlacement,
        max_value: Option<u64>,
        shape: HostShape,
        seed: HostSeed,
    ) -> Result<DeviceRing<T>> {
        let ctx = sess.b200().0;
        let dims = shape_i64(&shape);
        let mut out = std::ptr::null_mut();
        let mv = max_value.map(|v| v as i64).unwrap_or(-1);
        check(ctx, unsafe { mb_sample_seeded(ctx, T::DTYPE, dims.len() as i32, dims.as_ptr(), seed.0 .0.as_ptr(), mv, &mut out) })?;
        Ok(DeviceRing::from_raw(ctx, out, plc.clone()))
    }
    /// replaces moose/src/host/ops.rs:2019-2036
    pub(crate) fn bit_kernel<S: B200Session>(
        sess: &S,
        plc: &HostPlacement,
        max_value: Option<u64>,
        shape: HostShape,
        seed: HostSeed,
    ) -> Result<DeviceBits> {
        let ctx = sess.b200().0;
        let dims = shape_i64(&shape);
        let mut out = std::ptr::null_mut();
        let mv = max_value.map(|v| v as i64).unwrap_or(-1);
        check(ctx, unsafe { mb_sample_seeded(ctx, MB_BIT, dims.len() as i32, dims.as_ptr(), seed.0 .0.as_ptr(), mv, &mut out) })?;
        Ok(DeviceBits::from_raw(ctx, out, plc.clone()))
    }
}

impl crate::computation::SampleOp {
    /// replaces moose/src/host/ops.rs:1883-1957 (seed from `AesRng::from_random_seed`: the library draws it from the
    /// OS unless the context has a test entropy key)
    pub(crate) fn ring_kernel<S: B200Session, T: RingWidth>(
        sess: &S,
        plc: &HostPlacement,
        max_value: Option<u64>,
        shape: HostShape,
    ) -> Result<DeviceRing<T>> {
        let ctx = sess.b200().0;
        let dims = shape_i64(&shape);
        let mut out = std::ptr::null_mut();
        let mv = max_value.map(|v| v as i64).unwrap_or(-1);
        check(ctx, unsafe { mb_sample(ctx, T::DTYPE, dims.len() as i32, dims.as_ptr(), mv, &mut out) })?;
        Ok(DeviceRing::from_raw(ctx, out, plc.clone()))
    }
}

impl crate::computation::PrfKeyGenOp {
    /// replaces moose/src/host/prim.rs:113-121
    pub(crate) fn kernel<S: B200Session>(sess: &S, plc: &HostPlacement) -> Result<HostPrfKey> {
        let mut key = [0u8; 16];
        check(sess.b200().0, unsafe { mb_prf_key_gen(sess.b200().0, key.as_mut_ptr()) })?;
        Ok(HostPrfKey(RawPrfKey(key), plc.clone()))
    }
}

impl crate::computation::DeriveSeedOp {
    /// replaces moose/src/host/prim.rs:123-149: blake3 keyed hash of (session id || sync key); the session id is the
    /// context's (set from `sess.session_id()` when the context was created)
    pub(crate) fn kernel<S: B200Session>(sess: &S, plc: &HostPlacement, sync_key: SyncKey, key: HostPrfKey) -> Result<HostSeed> {
        let mut seed = [0u8; 16];
        check(sess.b200().0, unsafe { mb_derive_seed(sess.b200().0, sync_key.as_bytes().as_ptr(), key.0 .0.as_ptr(), seed.as_mut_ptr()) })?;
        Ok(HostSeed(RawSeed(seed), plc.clone()))
    }
}

impl crate::computation::ShlDimOp {
    /// replaces moose/src/host/ops.rs:835-853
    pub(crate) fn host_bit_kernel<S: B200Session>(
        sess: &S,
        plc: &HostPlacement,
        amount: usize,
        bit_length: usize,
        x: DeviceBits,
    ) -> Result<DeviceBits> {
        let ctx = sess.b200().0;
        let mut out = std::ptr::null_mut();
        check(ctx, unsafe { mb_shl_dim(ctx, amount as u64, bit_length as u64, x.raw(), &mut out) })?;
        Ok(DeviceBits::from_raw(ctx, out, plc.clone()))
    }
}

impl crate::computation::BitDecomposeOp {
    /// replaces moose/src/host/ops.rs:896-936 (host_bit64_kernel / host_bit128_kernel): [k, shape...] bit planes
    pub(crate) fn host_bit_kernel<S: B200Session, T: RingWidth>(sess: &S, plc: &HostPlacement, x: DeviceRing<T>) -> Result<DeviceBits> {
        let ctx = sess.b200().0;
        let mut out = std::ptr::null_mut();
        check(ctx, unsafe { mb_bit_decompose(ctx, x.raw(), &mut out) })?;
        Ok(DeviceBits::from_raw(ctx, out, plc.clone()))
    }
    /// replaces moose/src/host/ops.rs:856-894 (host_ring64_kernel / host_ring128_kernel): the planes as ring elements
    pub(crate) fn host_ring_kernel<S: B200Session, T: RingWidth>(sess: &S, plc: &HostPlacement, x: DeviceRing<T>) -> Result<DeviceRing<T>> {
        let ctx = sess.b200().0;
        let mut out = std::ptr::null_mut();
        check(ctx, unsafe { mb_bit_decompose_ring(ctx, x.raw(), &mut out) })?;
        Ok(DeviceRing::from_raw(ctx, out, plc.clone()))
    }
}

impl crate::computation::BitExtractOp {
    /// replaces moose/src/host/ops.rs:1567-1595 (kernel64 / kernel128)
    pub(crate) fn kernel<S: B200Session, T: RingWidth>(sess: &S, plc: &HostPlacement, bit_idx: usize, x: DeviceRing<T>) -> Result<DeviceBits> {
        let ctx = sess.b200().0;
        let mut out = std::ptr::null_mut();
        check(ctx, unsafe { mb_bit_extract(ctx, bit_idx as u64, x.raw(), &mut out) })?;
        Ok(DeviceBits::from_raw(ctx, out, plc.clone()))
    }
}

impl crate::computation::RingInjectOp {
    /// replaces moose/src/host/ops.rs:1597-1621
    pub(crate) fn host_kernel<S: B200Session, T: RingWidth>(sess: &S, plc: &HostPlacement, bit_idx: usize, x: DeviceBits) -> Result<DeviceRing<T>> {
        let ctx = sess.b200().0;
        let mut out = std::ptr::null_mut();
        check(ctx, unsafe { mb_ring_inject(ctx, T::DTYPE, bit_idx as u64, x.raw(), &mut out) })?;
        Ok(DeviceRing::from_raw(ctx, out, plc.clone()))
    }
}

macro_rules! bit_binary {
    ($op:ident, $ffi:ident, $cite:literal) => {
        impl crate::computation::$op {
            #[doc = $cite]
            pub(crate) fn host_kernel<S: B200Session>(sess: &S, plc: &HostPlacement, x: DeviceBits, y: DeviceBits) -> Result<DeviceBits> {
                let ctx = sess.b200().0;
                let mut out = std::ptr::null_mut();
                check(ctx, unsafe { $ffi(ctx, x.raw(), y.raw(), &mut out) })?;
                Ok(DeviceBits::from_raw(ctx, out, plc.clone()))
            }
        }
    };
}
bit_binary!(XorOp, mb_xor, "replaces moose/src/host/ops.rs:1507-1521");
bit_binary!(AndOp, mb_and, "replaces moose/src/host/ops.rs:1523-1537");
bit_binary!(OrOp, mb_or, "replaces moose/src/host/ops.rs:1539-1553");

impl crate::computation::AddNOp {
    /// replaces moose/src/host/ops.rs:1084-1103 (variadic)
    pub(crate) fn host_kernel<S: B200Session, T: RingWidth>(sess: &S, plc: &HostPlacement, xs: &[DeviceRing<T>]) -> Result<DeviceRing<T>> {
        if xs.is_empty() {
            return Err(Error::InvalidArgument("cannot reduce on empty array of tensors".to_string()));   // host/ops.rs:1090
        }
        let ctx = sess.b200().0;
        let raw: Vec<*const MbTensor> = xs.iter().map(|x| x.raw()).collect();
        let mut out = std::ptr::null_mut();
        check(ctx, unsafe { mb_add_n(ctx, raw.as_ptr(), raw.len() as i32, &mut out) })?;
        Ok(DeviceRing::from_raw(ctx, out, plc.clone()))
    }
}

impl crate::computation::SumOp {
    /// replaces moose/src/host/ops.rs:1067-1081
    pub(crate) fn host_ring_kernel<S: B200Session, T: RingWidth>(sess: &S, plc: &HostPlacement, axis: Option<usize>, x: DeviceRing<T>) -> Result<DeviceRing<T>> {
        let ctx = sess.b200().0;
        let mut out = std::ptr::null_mut();
        check(ctx, unsafe { mb_sum(ctx, x.raw(), axis.map(|a| a as i32).unwrap_or(-1), &mut out) })?;
        Ok(DeviceRing::from_raw(ctx, out, plc.clone()))
    }
}

impl crate::computation::FillOp {
    /// replaces moose/src/host/ops.rs:1635-1655 (ring64 / ring128; the constant arrives as the ring value,
    /// moose/src/kernels/constants.rs:59-130)
    pub(crate) fn host_ring_kernel<S: B200Session, T: RingWidth>(sess: &S, plc: &HostPlacement, value: u128, shape: HostShape) -> Result<DeviceRing<T>> {
        let ctx = sess.b200().0;
        let dims = shape_i64(&shape);
        let mut out = std::ptr::null_mut();
        check(ctx, unsafe { mb_fill(ctx, T::DTYPE, dims.len() as i32, dims.as_ptr(), value as u64, (value >> 64) as u64, &mut out) })?;
        Ok(DeviceRing::from_raw(ctx, out, plc.clone()))
    }
}

impl crate::computation::IndexAxisOp {
    /// replaces moose/src/host/ops.rs:740-783
    pub(crate) fn host_ring_kernel<S: B200Session, T: RingWidth>(sess: &S, plc: &HostPlacement, axis: usize, index: usize, x: DeviceRing<T>) -> Result<DeviceRing<T>> {
        let ctx = sess.b200().0;
        let mut out = std::ptr::null_mut();
        check(ctx, unsafe { mb_index_axis(ctx, axis as i32, index as i64, x.raw(), &mut out) })?;
        Ok(DeviceRing::from_raw(ctx, out, plc.clone()))
    }
}

impl crate::computation::ShapeOp {
    /// replaces moose/src/host/ops.rs:1657-1666: metadata only, no device work
    pub(crate) fn host_ring_kernel<S: B200Session, T: RingWidth>(_sess: &S, plc: &HostPlacement, x: DeviceRing<T>) -> Result<HostShape> {
        Ok(HostShape(RawShape(x.shape()), plc.clone()))
    }
}

// The remaining rows of the kernel table (Broadcast, Diag, ExpandDims, Squeeze, Concat, Transpose, Reshape, Select,
// Slice, Mux, Less, Greater, Sign, Cast, Maximum, RingFixedpointEncode / Decode / Mean / Argmax) follow the same
// three lines -- context, FFI call, re-stamp the placement -- over the entry points of the same names in b200_ffi.rs;
// the fused protocol entries (mb_rep_*) slot in one level up, as the bodies of the replicated-placement kernels of
// moose/src/replicated/*.rs, where a session runs all three parties in one process (LocalMooseRuntime).

// ---- registration -----------------------------------------------------------------------------------------------------------
// moose/src/kernels/arithmetic.rs:27-28 (and the other `modelled_kernel!` tables) gain one row per device type, e.g.
//
//     modelled_kernel! {
//         PlacementAdd::add, AddOp,
//         [
//             (HostPlacement, (HostRing64Tensor, HostRing64Tensor) -> HostRing64Tensor => [runtime] Self::ring_kernel),
//             #[cfg(feature = "b200")]
//             (HostPlacement, (DeviceRing64Tensor, DeviceRing64Tensor) -> DeviceRing64Tensor => [runtime] Self::ring_kernel),
//             ...
//
// and the lowering of Input / Constant / Load / Output / Save inserts `DeviceRing::from_host` / `to_host` at the edges.
