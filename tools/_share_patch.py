# scratch: applies the warp-local work sharing experiment to a copy of render_kernel.cuh (argv[1]); argv[2] = "eager" or "lazy"
import sys
k, mode = sys.argv[1], sys.argv[2]
s = open(k).read()
def sub(a, b):
    global s
    assert s.count(a) == 1, (s.count(a), a[:70])
    s = s.replace(a, b, 1)
sub('''enum { ST_IDLE = 0, ST_CLOSEST = 1, ST_CONTAINERS = 2, ST_SHADOW = 3 };''', '''enum { ST_IDLE = 0, ST_CLOSEST = 1, ST_CONTAINERS = 2, ST_SHADOW = 3, ST_JOIN = 4 };''')
sub('''    PK_SPARE = 15,   // (three free slots)''', '''    PK_HELPERS = 15,''')
sub('''    RayTask rays[RTC_RAY_STACK];
    int rsp = 0;
''', '''    RayTask rays[RTC_RAY_STACK];
    int rsp = 0;
    int owner = (int)threadIdx.x;
''')
sub('''                        pk_st(pk, PK_PCOL, v3(0, 0, 0));  // +0.0 is also the bit pattern of int64 0
                        rsp = 0;''', '''                        pk_st(pk, PK_PCOL, v3(0, 0, 0));  // +0.0 is also the bit pattern of int64 0
                        *reinterpret_cast<volatile int*>(&PK(PK_HELPERS)) = 0;
                        rsp = 0;''')
sub('''        if (__all_sync(0xffffffffu, done && state == ST_IDLE)) break;
''', '''        {
            unsigned takers = __ballot_sync(0xffffffffu, state == ST_IDLE);
            unsigned givers = __ballot_sync(0xffffffffu, rsp > 0 && state != ST_IDLE);
            if (takers && givers) {
                const int n = min(__popc(takers), __popc(givers));
                const bool give = rsp > 0 && state != ST_IDLE && __popc(givers & lt_mask) < n;
                const bool take = state == ST_IDLE && __popc(takers & lt_mask) < n;
                RayTask tk = rays[0];
                int tk_owner = owner;
                if (give) {
                    rays[0] = rays[rsp - 1];
                    rsp--;
                    atomicAdd(reinterpret_cast<int*>(s_mem + PK_HELPERS * RTC_BLOCK + owner), 1);
                }
                const int src = take ? (int)__fns(givers, 0, __popc(takers & lt_mask) + 1) : (int)lane;
                tk.ox = __shfl_sync(0xffffffffu, tk.ox, src);
                tk.oy = __shfl_sync(0xffffffffu, tk.oy, src);
                tk.oz = __shfl_sync(0xffffffffu, tk.oz, src);
                tk.dx = __shfl_sync(0xffffffffu, tk.dx, src);
                tk.dy = __shfl_sync(0xffffffffu, tk.dy, src);
                tk.dz = __shfl_sync(0xffffffffu, tk.dz, src);
                tk.w = __shfl_sync(0xffffffffu, tk.w, src);
                tk.remaining = __shfl_sync(0xffffffffu, tk.remaining, src);
                tk_owner = __shfl_sync(0xffffffffu, tk_owner, src);
                if (take) {
                    pk_st(pk, PK_RO, v3(tk.ox, tk.oy, tk.oz));
                    pk_st(pk, PK_RD, v3(tk.dx, tk.dy, tk.dz));
                    PK(PK_RW) = tk.w;
                    remaining = tk.remaining;
                    owner = tk_owner;
                    rsp = 0;
                    state = ST_CLOSEST;
                }
            }
        }
        if (__all_sync(0xffffffffu, done && state == ST_IDLE)) break;
''')
sub('''        if (state != ST_IDLE) {
            q.o = pk_ld(pk, PK_RO);''', '''        const bool tracing = state != ST_IDLE && state != ST_JOIN;
        if (tracing) {
            q.o = pk_ld(pk, PK_RO);''')
sub('''        busy = trace_begin<MESH>(q, S, trav, &err, state != ST_IDLE && !skip);''', '''        busy = trace_begin<MESH>(q, S, trav, &err, tracing && !skip);''')
sub('''                    long long a = __double_as_longlong(PK(PK_PCOL + k)) + __double2ll_rn(v[k]);
                    PK(PK_PCOL + k) = __longlong_as_double(a);''', '''                    long long a = __double2ll_rn(v[k]);
                    if (a) atomicAdd(reinterpret_cast<unsigned long long*>(s_mem + (PK_PCOL + k) * RTC_BLOCK + owner), (unsigned long long)a);''')
a = s.index('        if (pop_next) {\n            if (rsp > 0) {')
b = s.index('    // ---- ray counters: warp reduce')
s = s[:a] + '''        if (pop_next) {
            if (rsp > 0) {
                RayTask tk = rays[--rsp];
                pk_st(pk, PK_RO, v3(tk.ox, tk.oy, tk.oz));
                pk_st(pk, PK_RD, v3(tk.dx, tk.dy, tk.dz));
                PK(PK_RW) = tk.w;
                remaining = tk.remaining;
                state = ST_CLOSEST;
            } else if (owner != (int)threadIdx.x) {
                __threadfence_block();
                atomicSub(reinterpret_cast<int*>(s_mem + PK_HELPERS * RTC_BLOCK + owner), 1);
                owner = (int)threadIdx.x;
                state = ST_IDLE;
            } else {
                state = ST_JOIN;
            }
        }
        if (state == ST_JOIN && *reinterpret_cast<volatile int*>(&PK(PK_HELPERS)) == 0) {
            __threadfence_block();
            const volatile long long* acc = reinterpret_cast<const volatile long long*>(&PK(PK_PCOL));
            const double inv_scale = 1.0 / P.scale;  // a power of two: exact
            V3 pcol = v3((double)acc[0] * inv_scale, (double)acc[RTC_BLOCK] * inv_scale, (double)acc[2 * RTC_BLOCK] * inv_scale);
            if (P.out8) {
                uchar4 c = make_uchar4(quantise(pcol.x), quantise(pcol.y), quantise(pcol.z), 255);
                reinterpret_cast<uchar4*>(P.out8)[pix] = c;
            } else {
                double* o = P.out + pix * 3;
                o[0] = pcol.x;
                o[1] = pcol.y;
                o[2] = pcol.z;
            }
            state = ST_IDLE;
        }
    }

''' + s[b:]
if mode == "eager":
    sub('''                if (M.transparency != 0.0 && !S.ri_all_one) state = ST_CONTAINERS;
                else state = S.n_lights > 0 ? ST_SHADOW : -1;  // no lights: straight to the children below''', '''                if (M.transparency != 0.0 && !S.ri_all_one) {
                    state = ST_CONTAINERS;
                } else {
                    spawn = true;
                    state = S.n_lights > 0 ? ST_SHADOW : -1;
                }''')
    sub('''            state = S.n_lights > 0 ? ST_SHADOW : -1;
        } else if (state == ST_SHADOW) {''', '''            spawn = true;
            state = S.n_lights > 0 ? ST_SHADOW : -1;
        } else if (state == ST_SHADOW) {''')
    sub('''        bool pop_next = false;
        if (state == ST_CLOSEST) {''', '''        bool pop_next = false, spawn = false;
        if (state == ST_CLOSEST) {''')
    sub('''        if (state == -1) {
            // shadeHit tail''', '''        if (spawn) {
            // shadeHit tail''')
    sub('''            pop_next = true;
        }

        if (pop_next) {''', '''        }
        if (state == -1) pop_next = true;

        if (pop_next) {''')
open(k, 'w').write(s)
