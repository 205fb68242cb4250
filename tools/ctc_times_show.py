import csv, glob
for f in sorted(glob.glob('gpurun_out/ctc_l_*.csv')):
    rows = [r for r in csv.reader(open(f)) if len(r) > 10]
    h = rows[0]; kn = h.index('Kernel Name'); mn = h.index('Metric Name'); mv = h.index('Metric Value'); idi = h.index('ID')
    d = {}
    for r in rows[1:]:
        d.setdefault((r[idi], r[kn][:44]), {})[r[mn]] = r[mv]
    print(f)
    for k, v in d.items():
        if 'b2asr' in k[1] and 'rescale' not in k[1]:
            print('   %-46s %8.1f us  rd %7.1f MB  wr %7.1f MB' % (k[1], float(v['gpu__time_duration.sum']) / 1e3, float(v['dram__bytes_read.sum']) / 1e6, float(v['dram__bytes_write.sum']) / 1e6))
