python tools/gnn_convergence.py --n 7 --graphs 8 --steps 300 > gpurun_out/r2d_conv_7x7.json 2> gpurun_out/r2d_conv_7x7.err
python tools/gnn_convergence.py --n 20 --graphs 2 --steps 300 > gpurun_out/r2d_conv_20x20.json 2> gpurun_out/r2d_conv_20x20.err
python tools/gnn_convergence.py --n 7 --graphs 8 --steps 300 --lr 3e-4 > gpurun_out/r2d_conv_7x7_lr3e-4.json 2> gpurun_out/r2d_conv_7x7b.err
