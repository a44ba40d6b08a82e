/* Native row reader for GraphNeT SQLite databases (host-side input pipeline).
 *
 * Replaces the per-event Python queries of the reference's SQLiteDataset
 * (data/dataset/sqlite/sqlite_dataset.py:46-84: `conn.execute(...).fetchall()` -> np.asarray) with
 * one prepared statement stepped in C: the Python sqlite3 module spends ~0.4 us per fetched value
 * building tuples, which caps a reader at ~5 k events/s per thread; stepping the statement here and
 * writing doubles straight into the caller's buffer is an order of magnitude faster.
 *
 * Built by __graft_entry__.build() into libgn_sqlite.so against the system's libsqlite3.so.0 (the
 * few sqlite3 prototypes used are declared in sqlite_reader.c; the image ships no sqlite3.h).
 * Not part of the kernel ABI in include/graphnet_b200.h: plain host code, no CUDA. */
#ifndef GN_SQLITE_H
#define GN_SQLITE_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

/* Open read-only; NULL on failure (gn_sqlite_last_error() says why). */
void* gn_sqlite_open(const char* path);
void gn_sqlite_close(void* handle);
const char* gn_sqlite_last_error(void);

/* For every id in ids[0..n_ids): run `sql` - a SELECT of `ncols` numeric columns with exactly one
 * `?` parameter, the event number - and append its rows, as doubles, to out[rows][ncols] (row-major,
 * capacity cap_rows).  rows_per_id[i] receives the number of rows of ids[i]; rows of one event
 * come in the statement's order (rowid order for `WHERE event_no = ?`).
 * Returns the total number of rows, -1 on an SQLite error, -2 if cap_rows is too small (nothing
 * useful in out; call again with a larger buffer). */
int64_t gn_sqlite_read_events(void* handle, const char* sql, const int64_t* ids, int64_t n_ids,
                              int ncols, double* out, int64_t cap_rows, int64_t* rows_per_id);

#ifdef __cplusplus
}
#endif
#endif
