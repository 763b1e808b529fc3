/*
 * Arrow C Data Interface — the frozen, published ABI (arrow.apache.org/docs/format/CDataInterface.html).
 * Struct layouts are fixed by the Arrow specification; no Arrow library is required to use them.
 * This is how RecordBatches cross the dfgpu C ABI: the Rust side exports with `arrow::ffi`
 * (`FFI_ArrowArray` / `FFI_ArrowSchema` are these same structs), pyarrow with `_export_to_c`.
 */
#ifndef ARROW_C_DATA_INTERFACE
#define ARROW_C_DATA_INTERFACE

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ARROW_FLAG_DICTIONARY_ORDERED 1
#define ARROW_FLAG_NULLABLE 2
#define ARROW_FLAG_MAP_KEYS_SORTED 4

struct ArrowSchema {
    const char *format;
    const char *name;
    const char *metadata;
    int64_t flags;
    int64_t n_children;
    struct ArrowSchema **children;
    struct ArrowSchema *dictionary;
    void (*release)(struct ArrowSchema *);
    void *private_data;
};

struct ArrowArray {
    int64_t length;
    int64_t null_count;
    int64_t offset;
    int64_t n_buffers;
    int64_t n_children;
    const void **buffers;
    struct ArrowArray **children;
    struct ArrowArray *dictionary;
    void (*release)(struct ArrowArray *);
    void *private_data;
};

#ifdef __cplusplus
}
#endif
#endif /* ARROW_C_DATA_INTERFACE */
