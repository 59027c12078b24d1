import csv, re, sys, collections
sass_file, func, ncu_csv, rows_n = sys.argv[1], sys.argv[2], sys.argv[3], float(sys.argv[4])
lines=open(sass_file).read().split('\n')
start=[i for i,l in enumerate(lines) if l.startswith('.text.'+func+':')][0]
seq=[]  # (file,line) per instruction in order
cur=('?',0)
for l in lines[start+1:]:
    if l.startswith('//-----') : break
    m=re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m: cur=(m.group(1).split('/')[-1], int(m.group(2))); continue
    if re.match(r'\s+/\*[0-9a-f]{4,}\*/\s+\S', l): seq.append(cur)
rows=list(csv.reader(open(ncu_csv)))
hdr=rows[1]; ie=hdr.index('Instructions Executed')
data=rows[2:]
print(len(seq), len(data))
agg=collections.Counter()
for loc,r in zip(seq,data):
    agg[loc]+=int(r[ie])
tot=sum(agg.values())
print('total per row', tot/rows_n)
src={}
for (f,ln),c in agg.most_common(int(sys.argv[5]) if len(sys.argv)>5 else 40):
    if f not in src:
        try: src[f]=open('/root/repo/moss_tts_b200/csrc/'+f).read().split('\n')
        except Exception: src[f]=[]
    text=src[f][ln-1].strip()[:100] if 0<ln<=len(src[f]) else ''
    print(f"{c/rows_n:8.1f} {f}:{ln}  {text}")
