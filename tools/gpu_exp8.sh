cp hmp2g_b200/libdca_b200.so /tmp/default.so
cp tools/variants/diagcount.so hmp2g_b200/libdca_b200.so
timeout 300 python tools/diag_counts.py 2>&1 | tail -14
cp /tmp/default.so hmp2g_b200/libdca_b200.so
