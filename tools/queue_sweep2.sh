#!/bin/bash
# poll-interval sweep of the ticket queue (bench workload, chunk 64)
for s in ${@:-64 250 500 1000 2000}; do export SCG_QUEUE_SLEEP=$s; echo "sleep $s"; bash tools/queue_sweep.sh 64; done
