/^==/ {sec=$0} /^Problem at body/ {c[sec]++} END {for (k in c) print c[k], k}
