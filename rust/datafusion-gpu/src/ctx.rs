//! Safe handles over the C ABI: one `GpuCtx` per process and GPU (device, stream, memory pools, NCCL communicator) — the
//! `TaskContext` / `RuntimeEnv` analogue — and `DeviceBatch`, a `RecordBatch` resident in HBM.
use std::ffi::CString;
use std::os::raw::c_void;
use std::ptr;
use std::sync::Arc;

use arrow::array::{Array, StructArray};
use arrow::datatypes::{Schema, SchemaRef};
use arrow::ffi::{from_ffi, to_ffi, FFI_ArrowArray, FFI_ArrowSchema};
use arrow::record_batch::RecordBatch;
use datafusion_common::{DataFusionError, Result};

use crate::sys::*;

/// `dfgpu_options` with Rust-side defaults (`datafusion.execution.batch_size` etc. are filled in by the rule).
#[derive(Clone, Debug)]
pub struct GpuOptions {
    pub device: i32,
    pub batch_size: usize,
    /// bytes an operator may hold on the device; 0 = unlimited (`MemoryPool`, execution/src/memory_pool/mod.rs:83-108)
    pub memory_limit: i64,
    pub rank: i32,
    pub world: i32,
    /// 128-byte `ncclUniqueId` shared by all ranks (`GpuCtx::nccl_unique_id` on rank 0, distributed out of band)
    pub nccl_unique_id: Option<Vec<u8>>,
}

impl Default for GpuOptions {
    fn default() -> Self {
        Self { device: 0, batch_size: 8192, memory_limit: 0, rank: 0, world: 1, nccl_unique_id: None }
    }
}

pub struct GpuCtx {
    pub(crate) ptr: *mut dfgpu_ctx,
    pub options: GpuOptions,
    _uid: Option<Vec<u8>>, // the communicator id must outlive dfgpu_ctx_create only, kept for Debug
}
// The library orders all work of a context on its one stream and guards its pools with mutexes; operators created on
// a context may be driven from different tokio workers (one stream per partition), as `ExecutionPlan: Send + Sync` asks.
unsafe impl Send for GpuCtx {}
unsafe impl Sync for GpuCtx {}

impl GpuCtx {
    pub fn try_new(options: GpuOptions) -> Result<Arc<Self>> {
        let mut o: dfgpu_options = unsafe { std::mem::zeroed() };
        unsafe { dfgpu_options_default(&mut o) };
        o.device = options.device;
        o.batch_size = options.batch_size as i32;
        o.memory_limit = options.memory_limit;
        o.rank = options.rank;
        o.world = options.world;
        let uid = options.nccl_unique_id.clone();
        if let Some(u) = &uid {
            if u.len() != 128 {
                return Err(DataFusionError::Plan("ncclUniqueId must be 128 bytes".into()));
            }
            o.nccl_unique_id = u.as_ptr() as *const c_void;
        }
        let mut ptr = ptr::null_mut();
        check(unsafe { dfgpu_ctx_create(&o, &mut ptr) })?; // no CUDA device => DFGPU_CUDA: there is no CPU fallback
        Ok(Arc::new(Self { ptr, options, _uid: uid }))
    }

    /// rank 0 creates the communicator id and ships it to the other ranks
    pub fn nccl_unique_id() -> Result<Vec<u8>> {
        let mut id = vec![0u8; 128];
        check(unsafe { dfgpu_nccl_get_unique_id(id.as_mut_ptr() as *mut c_void) })?;
        Ok(id)
    }

    /// The row hash behind every partitioning decision of this library; see `dfgpu_hash_contract` in the header.
    pub fn hash_contract() -> i32 {
        unsafe { dfgpu_hash_contract() }
    }

    pub fn synchronize(&self) -> Result<()> {
        check(unsafe { dfgpu_ctx_synchronize(self.ptr) })
    }
}

impl Drop for GpuCtx {
    fn drop(&mut self) {
        unsafe { dfgpu_ctx_destroy(self.ptr) } // reference-counted inside: batches / operators may outlive this handle
    }
}

/// A `RecordBatch` in HBM (`dfgpu_batch`).
pub struct DeviceBatch {
    pub(crate) ptr: *mut dfgpu_batch,
    pub(crate) ctx: Arc<GpuCtx>,
}
unsafe impl Send for DeviceBatch {}

impl DeviceBatch {
    pub(crate) fn from_raw(ctx: &Arc<GpuCtx>, ptr: *mut dfgpu_batch) -> Option<Self> {
        if ptr.is_null() {
            None
        } else {
            Some(Self { ptr, ctx: ctx.clone() })
        }
    }

    /// host -> device.  The batch is borrowed for the duration of the call (Arrow C Data Interface consumer that
    /// never calls `release`); it may be dropped as soon as this returns.
    pub fn from_record_batch(ctx: &Arc<GpuCtx>, batch: &RecordBatch) -> Result<Self> {
        let (array, schema) = export_batch(batch)?;
        let mut out = ptr::null_mut();
        check(unsafe { dfgpu_batch_import(ctx.ptr, &array, &schema, 0, &mut out) })?;
        Ok(Self { ptr: out, ctx: ctx.clone() })
    }

    /// device -> host: library-allocated pinned buffers whose `release` callback returns them to the pool.
    pub fn to_record_batch(&self) -> Result<RecordBatch> {
        let mut array = FFI_ArrowArray::empty();
        let mut schema = FFI_ArrowSchema::empty();
        check(unsafe { dfgpu_batch_export(self.ctx.ptr, self.ptr, &mut array, &mut schema) })?;
        import_batch(array, &schema)
    }

    pub fn num_rows(&self) -> usize {
        unsafe { dfgpu_batch_num_rows(self.ptr) as usize }
    }

    /// `RecordBatch::project` on the device: the columns share the input's buffers (ProjectionExec of plain columns).
    pub fn project(&self, columns: &[usize]) -> Result<Self> {
        let cols: Vec<i32> = columns.iter().map(|c| *c as i32).collect();
        let mut out = ptr::null_mut();
        check(unsafe { dfgpu_batch_project(self.ctx.ptr, self.ptr, cols.as_ptr(), cols.len() as i32, &mut out) })?;
        Ok(Self { ptr: out, ctx: self.ctx.clone() })
    }
}

impl Drop for DeviceBatch {
    fn drop(&mut self) {
        unsafe { dfgpu_batch_free(self.ptr) }
    }
}

/// `RecordBatch -> StructArray -> FFI` (what pyarrow's `RecordBatch._export_to_c` produces).
pub(crate) fn export_batch(batch: &RecordBatch) -> Result<(FFI_ArrowArray, FFI_ArrowSchema)> {
    let s: StructArray = batch.clone().into();
    to_ffi(&s.to_data()).map_err(DataFusionError::from)
}

pub(crate) fn export_schema(schema: &Schema) -> Result<FFI_ArrowSchema> {
    FFI_ArrowSchema::try_from(schema).map_err(DataFusionError::from)
}

pub(crate) fn import_batch(array: FFI_ArrowArray, schema: &FFI_ArrowSchema) -> Result<RecordBatch> {
    let data = unsafe { from_ffi(array, schema) }?;
    Ok(RecordBatch::from(StructArray::from(data)))
}

pub(crate) fn import_schema(schema: &FFI_ArrowSchema) -> Result<SchemaRef> {
    Ok(Arc::new(Schema::try_from(schema)?))
}

pub(crate) fn cstring(s: &str) -> CString {
    CString::new(s.replace('\0', " ")).expect("no interior NUL")
}

/// Output slicing of `GroupedHashAggregateStream` / `HashJoinStream` (row_hash.rs:478-502): at most `batch_size` rows.
pub(crate) fn slice_batches(batch: RecordBatch, batch_size: usize) -> Vec<RecordBatch> {
    if batch.num_rows() <= batch_size {
        return vec![batch];
    }
    (0..batch.num_rows()).step_by(batch_size).map(|o| batch.slice(o, batch_size.min(batch.num_rows() - o))).collect()
}
