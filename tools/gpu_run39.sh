# Both bench arms once more on the final code (the drop-in record now has the view streams).
set -x
( time python bench.py ) > gpurun_out/r2z_bench.json 2> gpurun_out/r2z_bench.err; tail -4 gpurun_out/r2z_bench.err
( time python bench.py --impl reference ) > gpurun_out/r2z_ref.json 2> gpurun_out/r2z_ref.err; tail -4 gpurun_out/r2z_ref.err
