# ncu --set full of the LONGEST launch of one kernel of a command: a launch list first, then the
# capture of that launch.  Usage: bash tools/prof_longest.sh TAG KERNEL_REGEX -- command...
TAG=$1; KRE=$2; shift 3
"$@" > gpurun_out/${TAG}_plain.log 2>&1 || { tail -5 gpurun_out/${TAG}_plain.log; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:$KRE --csv --log-file gpurun_out/${TAG}_launches.csv "$@" > /dev/null 2>&1
IDX=$(python - <<PY
import csv
rows=[r for r in csv.reader(open("gpurun_out/${TAG}_launches.csv")) if len(r)>5]
h=rows[0]; vi=h.index("Metric Value"); ui=h.index("Metric Unit")
sc={"ns":1e-3,"us":1.0,"ms":1e3,"s":1e6}
t=[float(r[vi].replace(",",""))*sc.get(r[ui],1e-3) for r in rows[1:]]
k=max(range(len(t)), key=lambda i:t[i])
import sys
sys.stderr.write("launches %d, total %.1f us, longest #%d = %.1f us\n"%(len(t),sum(t),k,t[k]))
print(k)
PY
)
echo "capturing launch $IDX of $KRE"
ncu --set full --clock-control none --import-source on -k regex:$KRE -s $IDX -c 1 -f -o gpurun_out/${TAG} "$@" > gpurun_out/${TAG}_ncu.log 2>&1
tail -2 gpurun_out/${TAG}_ncu.log
