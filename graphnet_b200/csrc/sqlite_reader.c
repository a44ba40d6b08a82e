/* See gn_sqlite.h.  Plain C, linked against libsqlite3.so.0. */
#include "gn_sqlite.h"

#include <stdio.h>
#include <string.h>

/* the part of sqlite3's stable C API used here (no sqlite3.h in the image) */
typedef struct sqlite3 sqlite3;
typedef struct sqlite3_stmt sqlite3_stmt;
int sqlite3_open_v2(const char* filename, sqlite3** db, int flags, const char* vfs);
int sqlite3_close(sqlite3* db);
int sqlite3_prepare_v2(sqlite3* db, const char* sql, int nbyte, sqlite3_stmt** stmt,
                       const char** tail);
int sqlite3_bind_int64(sqlite3_stmt* stmt, int index, long long value);
int sqlite3_step(sqlite3_stmt* stmt);
double sqlite3_column_double(sqlite3_stmt* stmt, int col);
int sqlite3_column_count(sqlite3_stmt* stmt);
int sqlite3_reset(sqlite3_stmt* stmt);
int sqlite3_finalize(sqlite3_stmt* stmt);
const char* sqlite3_errmsg(sqlite3* db);
#define GN_SQLITE_OK 0
#define GN_SQLITE_ROW 100
#define GN_SQLITE_DONE 101
#define GN_SQLITE_OPEN_READONLY 0x00000001

static __thread char g_err[512];

const char* gn_sqlite_last_error(void) { return g_err; }

void* gn_sqlite_open(const char* path) {
  sqlite3* db = NULL;
  if (sqlite3_open_v2(path, &db, GN_SQLITE_OPEN_READONLY, NULL) != GN_SQLITE_OK) {
    snprintf(g_err, sizeof(g_err), "cannot open %s: %s", path, db ? sqlite3_errmsg(db) : "?");
    if (db) sqlite3_close(db);
    return NULL;
  }
  return db;
}

void gn_sqlite_close(void* handle) {
  if (handle) sqlite3_close((sqlite3*)handle);
}

int64_t gn_sqlite_read_events(void* handle, const char* sql, const int64_t* ids, int64_t n_ids,
                              int ncols, double* out, int64_t cap_rows, int64_t* rows_per_id) {
  sqlite3* db = (sqlite3*)handle;
  sqlite3_stmt* st = NULL;
  if (!db || !sql || !ids || !out || !rows_per_id || ncols < 1) {
    snprintf(g_err, sizeof(g_err), "gn_sqlite_read_events: bad arguments");
    return -1;
  }
  if (sqlite3_prepare_v2(db, sql, -1, &st, NULL) != GN_SQLITE_OK) {
    snprintf(g_err, sizeof(g_err), "prepare failed: %s", sqlite3_errmsg(db));
    return -1;
  }
  if (sqlite3_column_count(st) != ncols) {
    snprintf(g_err, sizeof(g_err), "statement yields %d columns, caller expects %d",
             sqlite3_column_count(st), ncols);
    sqlite3_finalize(st);
    return -1;
  }
  int64_t total = 0;
  for (int64_t i = 0; i < n_ids; ++i) {
    sqlite3_reset(st);
    sqlite3_bind_int64(st, 1, (long long)ids[i]);
    int64_t n = 0;
    int rc;
    while ((rc = sqlite3_step(st)) == GN_SQLITE_ROW) {
      if (total >= cap_rows) {
        sqlite3_finalize(st);
        snprintf(g_err, sizeof(g_err), "row buffer too small");
        return -2;
      }
      double* row = out + total * ncols;
      for (int c = 0; c < ncols; ++c) row[c] = sqlite3_column_double(st, c);
      ++total;
      ++n;
    }
    if (rc != GN_SQLITE_DONE) {
      snprintf(g_err, sizeof(g_err), "step failed: %s", sqlite3_errmsg(db));
      sqlite3_finalize(st);
      return -1;
    }
    rows_per_id[i] = n;
  }
  sqlite3_finalize(st);
  return total;
}
