python tools/ab_stage.py orb_slam2v2-1_b200/orbx/liborbx.so ab/*.so
for k in 1 3 4; do echo "ORBX_SPLIT=$k"; ORBX_SPLIT=$k python tools/ab_stage.py; done
