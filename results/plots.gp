# gnuplot script written by cuda_mpi_reductions_b200.utils.results.write_gnuplot
# data: <results>/<DTYPE>_<OP>.txt rows 'DTYPE OP RANKS GiB/s' (first line blank)
set terminal svg size 900,600 dynamic font 'Helvetica,13'
set datafile missing ''
set grid ytics lc rgb '#cccccc'
set logscale y 10
set xlabel 'GPUs (ranks)'
set ylabel 'bandwidth of the whole problem [GiB/s]'
set key outside right center box

set output 'int_reduce.svg'
set title 'INT: element-wise reduce to root over NVLink vs one-GPU scalar reduction'
plot \
  '/root/repo/results/INT_SUM.txt' using 3:4 with linespoints lw 2 pt 7 lc rgb '#2ca02c' title 'reduce SUM', \
  '/root/repo/results/INT_MIN.txt' using 3:4 with linespoints lw 2 pt 7 lc rgb '#1f77b4' title 'reduce MIN', \
  '/root/repo/results/INT_MAX.txt' using 3:4 with linespoints lw 2 pt 7 lc rgb '#d62728' title 'reduce MAX', \
  3876.825 with lines dt 2 lw 2 lc rgb '#2ca02c' title 'one GPU SUM', \
  3881.023 with lines dt 2 lw 2 lc rgb '#1f77b4' title 'one GPU MIN', \
  3892.471 with lines dt 2 lw 2 lc rgb '#d62728' title 'one GPU MAX'

set output 'double_reduce.svg'
set title 'DOUBLE: element-wise reduce to root over NVLink vs one-GPU scalar reduction'
plot \
  '/root/repo/results/DOUBLE_SUM.txt' using 3:4 with linespoints lw 2 pt 7 lc rgb '#2ca02c' title 'reduce SUM', \
  '/root/repo/results/DOUBLE_MIN.txt' using 3:4 with linespoints lw 2 pt 7 lc rgb '#1f77b4' title 'reduce MIN', \
  '/root/repo/results/DOUBLE_MAX.txt' using 3:4 with linespoints lw 2 pt 7 lc rgb '#d62728' title 'reduce MAX', \
  4802.314 with lines dt 2 lw 2 lc rgb '#2ca02c' title 'one GPU SUM', \
  4773.645 with lines dt 2 lw 2 lc rgb '#1f77b4' title 'one GPU MIN', \
  4769.973 with lines dt 2 lw 2 lc rgb '#d62728' title 'one GPU MAX'
