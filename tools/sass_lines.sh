#!/bin/bash
# Static SASS instruction count per source line of one kernel of an object file:  tools/sass_lines.sh obj.o kernel_substring [N]
set -e
d=$(mktemp -d); cd $d
cuobjdump -xelf all "$1" > /dev/null
nvdisasm -g *.cubin 2>/dev/null > all.txt
awk -v k="$2" '/\.section[ \t]+\.text\./{insec = (index($0, k) > 0)} insec' all.txt | grep -E "//## File|^[ \t]+/\*[0-9a-f]{4,}\*/" | \
  awk '/## File/{match($0,/"[^"]*"/); f=substr($0,RSTART+1,RLENGTH-2); n=split(f,a,"/"); match($0,/line [0-9]+/); l=substr($0,RSTART+5,RLENGTH-5); cur=a[n]":"l; next} {cnt[cur]++; tot++} END{for(k in cnt) print cnt[k],k; print tot, "TOTAL"}' | sort -n -r | head -${3:-40}
rm -rf $d
