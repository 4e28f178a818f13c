//! never compiled here (no cargo in this image). ballista-gpu: the drop-in of the aggregation + shuffle hot path.
//! Registration, as the reference's own integration tests do it:
//!   scheduler: `new_standalone_scheduler_with_codec(BallistaCodec::new(logical, Arc::new(GpuPhysicalCodec::default())))`
//!              (/root/reference/ballista/scheduler/src/standalone.rs:42-45)
//!   executor:  `ExecutorProcessConfig { execution_engine: Some(Arc::new(GpuExecutionEngine { device, fallback })), .. }`
//!              (/root/reference/ballista/executor/src/executor_process.rs:92-94)
pub mod codec;
pub mod engine;
pub mod ffi;
pub mod gpu_aggregate;
pub mod gpu_shuffle_writer;
