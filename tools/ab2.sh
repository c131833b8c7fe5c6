mkdir -p gpurun_out
for v in base fair base fair; do
PTB_LIB=$PWD/build/ab/libptb_$v.so PTB_LAUNCH_LOG=gpurun_out/launchlog_$v.txt python tools/iterlog.py c3 64 | grep timing | sed "s/^/$v /"
done
