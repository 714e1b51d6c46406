import csv, sys, collections
rows=[r for r in csv.reader(open(sys.argv[1])) if len(r)>10]
hdr=rows[0]; ki=hdr.index("Kernel Name"); vi=hdr.index("Metric Value")
names=[r[ki].split("(")[0][-44:] for r in rows[1:]]; vals=[float(r[vi])/1e3 for r in rows[1:]]
n=int(sys.argv[2])
for i in range(max(0,len(names)-n), len(names)): print("%-46s %8.1f us" % (names[i], vals[i]))
