#!/usr/bin/env bash
# Build the TensorFlow custom-op shim against an installed TensorFlow (not available in this repository's image).
set -euo pipefail
HERE="$(cd "$(dirname "$0")" && pwd)"
python -c "import tensorflow" 2>/dev/null || { echo "TensorFlow is not installed: the shim cannot be built here."; exit 3; }
TF_CFLAGS=( $(python -c 'import tensorflow as tf; print(" ".join(tf.sysconfig.get_compile_flags()))') )
TF_LFLAGS=( $(python -c 'import tensorflow as tf; print(" ".join(tf.sysconfig.get_link_flags()))') )
python -m gcnn_keras_b200.build
g++ -std=c++17 -shared -fPIC -O2 "$HERE/kgcnn_tf_ops.cc" -o "$HERE/libkgcnn_tf_ops.so" \
    "${TF_CFLAGS[@]}" "${TF_LFLAGS[@]}" -DGOOGLE_CUDA=1 -I/usr/local/cuda/include \
    -L"$HERE/../gcnn_keras_b200" -lkgcnn_b200 -Wl,-rpath,"$HERE/../gcnn_keras_b200"
echo "built $HERE/libkgcnn_tf_ops.so"
