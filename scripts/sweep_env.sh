# usage: sweep_env.sh "VAR=a VAR2=b" "VAR=c" ...   -> median decode ms/step of profile_path.py under each env set
for envs in "$@"; do
  echo "== $envs"; env $envs python scripts/profile_path.py --what lm --steps ${STEPS:-60} 2>&1 | python -c "
import sys
for l in sys.stdin:
    if l.startswith('lm:'):
        v=[float(x) for x in l.split('decode steps')[1].split()]
        t=sorted(v[20:]); print('median %.3f min %.3f' % (t[len(t)//2], t[0]))
    elif 'rror' in l: print(l.strip())
"
done
