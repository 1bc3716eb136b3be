mkdir -p gpurun_out
timeout 1500 compute-sanitizer --tool memcheck --error-exitcode 7 python -m pytest tests/test_gpu_mc.py "tests/test_gpu_streams.py::test_stream_md5" tests/test_gpu_frames.py -m gpu -x -q -k "not 2160p and not 1080p and not concurrent" > gpurun_out/sanitizer.log 2>&1
echo "exit $?"
grep -c "Invalid\|out of bounds" gpurun_out/sanitizer.log
tail -5 gpurun_out/sanitizer.log
