#!/bin/sh
# Builds libduckdb.so (DuckDB v1.3.2) from the DuckDB sources of the reference tree, out of tree, and installs it
# as ext/_duckdb/libduckdb.so for ext/duckdb_sql_check (tests/test_duckdb_sql.py).  Dev container only (needs
# /root/reference/duckdb and about 6 minutes on 8 cores); `python __graft_entry__.py` then links the check driver.
# Nothing of DuckDB is copied into the repository: the library is a git-ignored build artefact.
set -e
HERE=$(cd "$(dirname "$0")" && pwd)
SRC=${1:-/root/reference/duckdb}
B=${2:-/tmp/duckdb_build}
cmake -G Ninja -DCMAKE_BUILD_TYPE=Release -DOVERRIDE_GIT_DESCRIBE=v1.3.2 -DENABLE_SANITIZER=FALSE -DENABLE_UBSAN=0 \
      -DBUILD_UNITTESTS=0 -DBUILD_SHELL=0 -DBUILD_EXTENSIONS="" -DDISABLE_BUILTIN_EXTENSIONS=1 -S "$SRC" -B "$B"
ninja -C "$B" duckdb
mkdir -p "$HERE/_duckdb"
cp "$B/src/libduckdb.so" "$HERE/_duckdb/libduckdb.so"
strip "$HERE/_duckdb/libduckdb.so"
