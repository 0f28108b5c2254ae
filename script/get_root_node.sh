#!/bin/bash
# Print the scheduler's node text for this machine:  script/get_root_node.sh <interface|-> <port>
# "-" (or an interface without an address) falls back to the first non-loop-back IPv4 address,
# then to 127.0.0.1.
if [ $# -ne 2 ]; then
  echo "usage: $0 interface port"
  exit 1
fi
ip=""
if [ "$1" != "-" ] && command -v ip >/dev/null 2>&1; then
  ip=$(ip -4 -o addr show dev "$1" 2>/dev/null | awk '{print $4}' | cut -d/ -f1 | head -n1)
fi
if [ -z "${ip}" ]; then
  ip=$(hostname -I 2>/dev/null | tr ' ' '\n' | grep -v '^127\.' | grep -v ':' | head -n1)
fi
if [ -z "${ip}" ]; then ip=127.0.0.1; fi
echo "role:SCHEDULER,hostname:'${ip}',port:$2,id:'H'"
