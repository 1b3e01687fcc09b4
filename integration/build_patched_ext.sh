#!/bin/sh
# Build the reference's SQLite extension with its hnsw_index KNN plan routed through
# libmuninn_cuda: the reference sources are compiled from where they lie under $REF; only
# hnsw_vtab.c is edited, in a scratch copy under $TMP, by the seven one-line substitutions
# below (the same edits INTEGRATION.md lists for a maintainer). muninn.c is unchanged.
# Output: integration/_build/muninn_cuda_ext.so (git-ignored; travels to the GPU box).
set -eu
HERE=$(cd "$(dirname "$0")" && pwd)
ROOT=$(dirname "$HERE")
REF=${REF:-/root/reference}
[ -f "$REF/src/hnsw_vtab.c" ] || { echo "integration: $REF absent — keeping prebuilt integration/_build if present" >&2; exit 0; }
CC=${HOSTCC:-/usr/bin/gcc}
[ -x "$CC" ] || CC=gcc
OUT="$HERE/_build/muninn_cuda_ext.so"
if [ -f "$OUT" ]; then
    fresh=1
    for f in "$HERE/muninn_vtab_glue.c" "$HERE/muninn_batch_tvf.c" "$HERE/muninn_vtab_glue.h" "$0" "$ROOT/include/muninn_cuda.h" "$REF/src/hnsw_vtab.c" "$REF/src/hnsw_algo.c"; do
        [ "$OUT" -nt "$f" ] || fresh=0
    done
    [ $fresh = 1 ] && exit 0
fi
TMP=$(mktemp -d)
trap 'rm -rf "$TMP"' EXIT
sed -e 's|cur->result_count = hnsw_search(vtab->index, query, k, ef_search, cur->results);|cur->result_count = muninn_glue_search(vtab->index, query, k, ef_search, cur->results, \&vtab->base.zErrMsg); if (vtab->base.zErrMsg) return SQLITE_ERROR;|' \
    -e 's|rc = hnsw_delete(vtab->index, id);|rc = hnsw_delete(vtab->index, id); if (rc == 0) muninn_glue_note_delete(vtab->index, id);|' \
    -e 's|rc = hnsw_insert(vtab->index, id, vec);|rc = hnsw_insert(vtab->index, id, vec); if (rc == 0) muninn_glue_note_insert(vtab->index, id);|' \
    -e '0,/int expected_bytes = /s|^\( *\)int expected_bytes = vtab->dim \* (int)sizeof(float);|\1if (idxNum == PLAN_KNN \&\& muninn_glue_batch_request(vtab->index, vtab->dim, argv[0], k, ef_search, \&vtab->base.zErrMsg)) return vtab->base.zErrMsg ? SQLITE_ERROR : SQLITE_OK; int expected_bytes = vtab->dim * (int)sizeof(float);|' \
    -e 's|return sqlite3_create_module(db, "hnsw_index", \&hnsw_module, NULL);|{ int rc_ = sqlite3_create_module(db, "hnsw_index", \&hnsw_module, NULL); return rc_ != SQLITE_OK ? rc_ : muninn_glue_register_batch_tvf(db); }|' \
    -e 's|^\( *\)hnsw_destroy(vtab->index);|\1muninn_glue_release(vtab->index); hnsw_destroy(vtab->index);|' \
    "$REF/src/hnsw_vtab.c" > "$TMP/hnsw_vtab.c"
grep -q muninn_glue_search "$TMP/hnsw_vtab.c" || { echo "integration: call site not found in hnsw_vtab.c" >&2; exit 1; }
[ "$(grep -c "muninn_glue_note_" "$TMP/hnsw_vtab.c")" = 2 ] || { echo "integration: xUpdate anchors not found" >&2; exit 1; }
[ "$(grep -c "muninn_glue_batch_request\|muninn_glue_register_batch_tvf" "$TMP/hnsw_vtab.c")" = 2 ] || { echo "integration: batch anchors not found" >&2; exit 1; }
[ "$(grep -c muninn_glue_release "$TMP/hnsw_vtab.c")" -ge 1 ] || { echo "integration: hnsw_destroy anchors not found" >&2; exit 1; }
SRCS=""
for f in "$REF"/src/*.c; do
    case "$(basename "$f")" in
        llama_*|sqlite3_wasm_extra_init.c|hnsw_vtab.c) ;;
        *) SRCS="$SRCS $f" ;;
    esac
done
mkdir -p "$HERE/_build"
CSRC="$ROOT/sqlite_muninn_b200/csrc"
$CC -O2 -std=c11 -fPIC -D_POSIX_C_SOURCE=200809L -DMUNINN_NO_LLAMA -shared -Wl,-Bsymbolic \
    -I"$REF/src" -I"$REF/vendor/yyjson" -I"$ROOT/include" -I"$HERE" \
    -include "$HERE/muninn_vtab_glue.h" \
    -o "$HERE/_build/muninn_cuda_ext.so" \
    "$TMP/hnsw_vtab.c" "$HERE/muninn_vtab_glue.c" "$HERE/muninn_batch_tvf.c" $SRCS "$REF"/vendor/yyjson/*.c \
    -L"$CSRC" -lmuninn_cuda -Wl,-rpath,'$ORIGIN/../../sqlite_muninn_b200/csrc' -lm
echo "integration: built $HERE/_build/muninn_cuda_ext.so" >&2
