/*
 * sqlite3_abi.h -- the handful of SQLite C-API entry points the FastAAI reader needs, declared by hand.
 *
 * The image ships the SQLite runtime (/usr/lib/x86_64-linux-gnu/libsqlite3.so.0, 3.45.1) but no
 * development header, and the reference's vendored header may not be copied.  The SQLite C API is a
 * stable public ABI (https://sqlite.org/capi3ref.html); the prototypes and result codes below are
 * written from that documentation.  Link with the runtime library directly.
 */
#ifndef PFAAI_SQLITE3_ABI_H
#define PFAAI_SQLITE3_ABI_H

#ifdef __cplusplus
extern "C" {
#endif

typedef struct sqlite3 sqlite3;
typedef struct sqlite3_stmt sqlite3_stmt;
typedef long long sqlite3_int64;

#define SQLITE_OK 0
#define SQLITE_ERROR 1
#define SQLITE_CANTOPEN 14
#define SQLITE_ROW 100
#define SQLITE_DONE 101

#define SQLITE_OPEN_READONLY 0x00000001
#define SQLITE_OPEN_NOMUTEX 0x00008000

int sqlite3_open(const char *filename, sqlite3 **db);
int sqlite3_open_v2(const char *filename, sqlite3 **db, int flags, const char *vfs);
int sqlite3_close(sqlite3 *db);
int sqlite3_exec(sqlite3 *db, const char *sql, int (*cb)(void *, int, char **, char **), void *arg, char **errmsg);
int sqlite3_prepare_v2(sqlite3 *db, const char *sql, int nbyte, sqlite3_stmt **stmt, const char **tail);
int sqlite3_step(sqlite3_stmt *stmt);
int sqlite3_finalize(sqlite3_stmt *stmt);
int sqlite3_column_int(sqlite3_stmt *stmt, int col);
sqlite3_int64 sqlite3_column_int64(sqlite3_stmt *stmt, int col);
const unsigned char *sqlite3_column_text(sqlite3_stmt *stmt, int col);
const void *sqlite3_column_blob(sqlite3_stmt *stmt, int col);
int sqlite3_column_bytes(sqlite3_stmt *stmt, int col);
const char *sqlite3_errmsg(sqlite3 *db);
const char *sqlite3_errstr(int rc);
const char *sqlite3_libversion(void);

#ifdef __cplusplus
}
#endif
#endif /* PFAAI_SQLITE3_ABI_H */
