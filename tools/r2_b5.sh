for w in 1.65 0; do echo "== trace w=$w"; FINDB200_DW_SPLIT=$w timeout 200 python tools/dw_trace.py 2>&1 | grep -A36 "^latest" | grep -v "cloud_sort" ; done
