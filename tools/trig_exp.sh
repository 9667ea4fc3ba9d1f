for cfg in "1 1 1" "1 0 1" "0 0 0" "1 0 0" "0 0 1" "0 1 0"; do set -- $cfg
QUEST_EXTRA_FLAGS="-DQUEST_TRIG_EST=$1 -DQUEST_TRIG_TOPK=$2 -DQUEST_TRIG_DEC=$3" bash quest_b200/csrc/build.sh > /dev/null 2>&1
echo "== EST=$1 TOPK=$2 DEC=$3"; timeout 200 python tools/chain_probe.py 2>&1 | grep -E "^estimate\+topk |^topk\+decode|^append\+estimate\+topk\+decode|^estimate\+topk\+decode|^estimate  |^decode  "
done
