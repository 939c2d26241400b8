import re,sys,collections
acc=collections.OrderedDict()
for l in sys.stdin:
    if 'gemm dbg' not in l: continue
    key=l.split('|')[0].strip()
    nums=[int(x) for x in re.findall(r'(?<![A-Za-z=])(\d+)(?= cyc|,|\)|\]| \[| \|)', l.split('|',1)[1])]
    acc.setdefault(key,[]).append(l.split('|',1)[1].strip())
for k,v in acc.items():
    print(k, 'x%d'%len(v)); print('   ', v[len(v)//2][:420])
