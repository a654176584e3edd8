#!/bin/sh
# oracle/cc_ref.sh OBJDIR REFROOT SRC -- [cc flags...]
# Compiles ONE reference source where it lies under REFROOT into OBJDIR (test infrastructure).
objdir="$1"; ref="$2"; src="$3"; shift 4
name=$(printf '%s' "${src#$ref/}" | tr '/' '_')
exec ${CC:-gcc} "$@" -c "$src" -o "$objdir/${name%.c}.o"
