cd metal-renderer_b200
for sc in spheres-1m cornell-water many-lights; do
for ml in 3 1; do
echo "== $sc maxleaf $ml"
MRT_BVH8_MAXLEAF=$ml MRT_TRACE_COUNT=1 ./bin/mrt_render --scene $sc --size 1920x1080 --seed 3 --frames 8 --pipeline wavefront | python -c "import sys,json; d=json.load(sys.stdin); r=d['primary_rays']+d['shadow_rays']; print('nodes/ray %.2f tris/ray %.2f bvh8_nodes %d depth %d'%(d['node_tests']/r, d['triangle_tests']/r, d['bvh8_nodes'], d['bvh8_depth']))"
for i in 1 2 3; do MRT_BVH8_MAXLEAF=$ml ./bin/mrt_render --scene $sc --size 1920x1080 --seed 3 --frames 64 --pipeline wavefront | python -c "import sys,json; d=json.load(sys.stdin); print('  device_ms %.2f mrays %.1f'%(d['device_ms'], d['mrays_per_s']))"; done
done; done
