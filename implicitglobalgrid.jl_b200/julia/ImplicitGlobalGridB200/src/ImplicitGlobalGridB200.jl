# ImplicitGlobalGridB200.jl -- Julia host side of libigg_b200.so.
#
# UNVERIFIED: Julia and MPI are not available in the environment this repository is built and tested in; this file
# has never been executed.  It is the binding a Julia user would load instead of ImplicitGlobalGrid.jl to get the
# same API for the halo path (init_global_grid, update_halo!, gather!, finalize_global_grid, select_device,
# nx_g/x_g/ix_g ..., tic/toc) on an 8xB200 box.  The Python twin (implicitglobalgrid.jl_b200/igg_b200), which makes
# the same C calls in the same order, is what the test-suite exercises.
#
# Division of labour (see INTEGRATION.md): MPI.jl provides the process topology and the one-time all-gather of the
# bootstrap blobs; everything that moves field data is one ccall into libigg_b200.so.
module ImplicitGlobalGridB200

using CUDA
import MPI

export init_global_grid, finalize_global_grid, update_halo!, gather!, select_device, get_global_grid
export nx_g, ny_g, nz_g, x_g, y_g, z_g, ix_g, iy_g, iz_g, tic, toc
export prepare_update_halo, alloc_field, free_field

const LIB = get(ENV, "IGG_B200_LIB", "libigg_b200.so")
const NDIMS = 3

# == igg_grid_desc (include/igg_b200.h)
struct GridDesc
    nxyz::NTuple{3,Int64}
    dims::NTuple{3,Int64}
    coords::NTuple{3,Int64}
    neighbors::NTuple{6,Int64}     # [d][side]: (left_x, right_x, left_y, right_y, left_z, right_z)
    periods::NTuple{3,Int64}
    overlaps::NTuple{3,Int64}
    halowidths::NTuple{3,Int64}
    me::Int32
    nprocs::Int32
    device::Int32
    reserved::Int32
end

# == igg_field
struct Field
    ptr::Ptr{Cvoid}
    size::NTuple{3,Int64}
    halowidths::NTuple{3,Int64}
    elsize::Int32
    ndims::Int32
end

mutable struct GlobalGrid
    nxyz_g::Vector{Int}
    nxyz::Vector{Int}
    dims::Vector{Int}
    overlaps::Vector{Int}
    halowidths::Vector{Int}
    origin::Vector{Float64}
    origin_on_vertex::Bool
    centerxyz::Vector{Bool}
    nprocs::Int
    me::Int
    coords::Vector{Int}
    neighbors::Matrix{Int}          # neighbors[side, dim] as in the reference
    periods::Vector{Int}
    comm::MPI.Comm
    quiet::Bool
    ctx::Ptr{Cvoid}
end

const GRID = Ref{Union{GlobalGrid,Nothing}}(nothing)
const T0 = Ref(0.0)

lasterror() = unsafe_string(ccall((:igg_last_error, LIB), Cstring, ()))
check(rc) = rc == 0 ? nothing : error(lasterror())          # every failure is an ErrorException, like the reference
grid() = (GRID[] === nothing ? error("No function of the module can be called before init_global_grid().") : GRID[])
get_global_grid() = deepcopy(grid())

"""
    init_global_grid(nx, ny, nz; dimx=0, ..., periodx=0, ..., overlaps=(2,2,2), halowidths=max.(1, overlaps .÷ 2), ...)

Same arguments and return values `(me, dims, nprocs, coords, comm_cart)` as ImplicitGlobalGrid.init_global_grid.
"""
function init_global_grid(nx::Integer, ny::Integer=1, nz::Integer=1; dimx::Integer=0, dimy::Integer=0, dimz::Integer=0,
                          periodx::Integer=0, periody::Integer=0, periodz::Integer=0, origin=(0.0, 0.0, 0.0),
                          origin_on_vertex::Bool=false, centerx::Bool=false, centery::Bool=false, centerz::Bool=false,
                          overlaps::NTuple{3,Int}=(2, 2, 2), halowidths::NTuple{3,Int}=max.(1, overlaps .÷ 2),
                          disp::Integer=1, reorder::Integer=1, comm::MPI.Comm=MPI.COMM_WORLD, init_MPI::Bool=true,
                          select_device::Bool=true, quiet::Bool=false)
    GRID[] === nothing || error("init_global_grid() can only be called once before finalize_global_grid().")
    nxyz = [nx, ny, nz]; dims = [dimx, dimy, dimz]; periods = [periodx, periody, periodz]
    any(dims .< 0) && error("Invalid arguments: dimx, dimy, and dimz cannot be negative.")
    all(p -> p in (0, 1), periods) || error("Invalid arguments: periodx, periody, and periodz must be either 0 or 1.")
    any(halowidths .< 1) && error("Invalid arguments: halowidths cannot be less than 1.")
    any((overlaps .> 0) .& (halowidths .> overlaps .÷ 2)) && error("Incoherent arguments: if overlap is greater than 0, then halowidth cannot be greater than overlap÷2, in each dimension.")
    any(nxyz .< 1) && error("Invalid arguments: nx, ny, and nz cannot be less than 1.")
    nx == 1 && error("Invalid arguments: nx can never be 1.")
    (ny == 1 && nz > 1) && error("Invalid arguments: ny cannot be 1 if nz is greater than 1.")
    any((nxyz .== 1) .& (dims .> 1)) && error("Incoherent arguments: if nx, ny, or nz is 1, then the corresponding dimx, dimy or dimz must not be set (or set 0 or 1).")
    any((nxyz .< 2 .* overlaps .- 1) .& (periods .> 0)) && error("Incoherent arguments: if nx, ny, or nz is smaller than 2*overlaps[1]-1, 2*overlaps[2]-1 or 2*overlaps[3]-1, respectively, then the corresponding periodx, periody or periodz must not be set (or set 0).")
    disp == 1 || error("libigg_b200 supports disp=1 only.")
    dims[(nxyz .== 1) .& (dims .== 0)] .= 1
    if init_MPI
        MPI.Initialized() && error("MPI is already initialized. Set the argument 'init_MPI=false'.")
        MPI.Init()
    else
        MPI.Initialized() || error("MPI has not been initialized beforehand. Remove the argument 'init_MPI=false'.")
    end
    nprocs = MPI.Comm_size(comm)
    dims .= MPI.Dims_create(nprocs, dims)
    comm_cart = MPI.Cart_create(comm, dims; periodic=periods .> 0, reorder=reorder > 0)
    me = MPI.Comm_rank(comm_cart)
    coords = MPI.Cart_coords(comm_cart)
    neighbors = fill(Int(MPI.PROC_NULL), 2, NDIMS)
    for d in 1:NDIMS
        neighbors[:, d] .= MPI.Cart_shift(comm_cart, d - 1, disp)
    end
    nxyz_g = dims .* (nxyz .- collect(overlaps)) .+ collect(overlaps) .* (periods .== 0)
    org = Float64[(origin isa Number ? (origin, 0.0, 0.0) : (origin..., 0.0, 0.0)[1:3])...]
    gg = GlobalGrid(nxyz_g, nxyz, dims, collect(overlaps), collect(halowidths), org, origin_on_vertex,
                    [centerx, centery, centerz], nprocs, me, coords, neighbors, periods, comm_cart, quiet, C_NULL)
    GRID[] = gg            # select_device() and tic()/toc() below read it; reset if anything fails (a retry must work)
    try
    (!quiet && me == 0) && println("Global grid: $(nxyz_g[1])x$(nxyz_g[2])x$(nxyz_g[3]) (nprocs: $nprocs, dims: $(dims[1])x$(dims[2])x$(dims[3]); device support: CUDA (libigg_b200))")
    dev = select_device ? ImplicitGlobalGridB200.select_device() : CUDA.deviceid(CUDA.device())
    nb = ntuple(i -> (v = neighbors[(i - 1) % 2 + 1, (i - 1) ÷ 2 + 1]; Int64(v == MPI.PROC_NULL ? -1 : v)), 6)
    desc = Ref(GridDesc(Tuple(nxyz), Tuple(dims), Tuple(coords), nb, Tuple(periods), overlaps, halowidths,
                        me, nprocs, dev, 0))
    ctx = Ref{Ptr{Cvoid}}(C_NULL)
    check(ccall((:igg_create, LIB), Cint, (Ref{GridDesc}, Ref{Ptr{Cvoid}}), desc, ctx))
    gg.ctx = ctx[]
    if nprocs > 1      # one-time rank bootstrap: the only MPI traffic of the halo path
        n = ccall((:igg_bootstrap_size, LIB), Csize_t, ())
        blob = Vector{UInt8}(undef, n)
        check(ccall((:igg_bootstrap_export, LIB), Cint, (Ptr{Cvoid}, Ptr{UInt8}), gg.ctx, blob))
        blobs = MPI.Allgather(blob, comm_cart)
        check(ccall((:igg_bootstrap_import, LIB), Cint, (Ptr{Cvoid}, Ptr{UInt8}, Cint), gg.ctx, blobs, nprocs))
    end
    tic(); toc()
    catch
        gg.ctx == C_NULL || ccall((:igg_destroy, LIB), Cint, (Ptr{Cvoid},), gg.ctx)
        GRID[] = nothing
        rethrow()
    end
    return me, dims, nprocs, coords, comm_cart
end

"Select the GPU with the node-local rank as device id (same rule as the reference)."
function select_device()
    gg = grid()
    comm_l = MPI.Comm_split_type(gg.comm, MPI.COMM_TYPE_SHARED, gg.me)
    MPI.Comm_size(comm_l) > length(CUDA.devices()) && error("More processes have been launched per node than there are GPUs available.")
    me_l = MPI.Comm_rank(comm_l)
    CUDA.device!(me_l)
    return me_l
end

function finalize_global_grid(; finalize_MPI::Bool=true)
    gg = grid()
    gg.ctx == C_NULL || check(ccall((:igg_destroy, LIB), Cint, (Ptr{Cvoid},), gg.ctx))
    if finalize_MPI
        MPI.Finalized() && error("MPI is already finalized. Set the argument 'finalize_MPI=false'.")
        MPI.Finalize()
    end
    GRID[] = nothing
    return nothing
end

# ---- update_halo! ------------------------------------------------------------------------------------------------
const FieldArg = Union{AbstractArray,NamedTuple{(:A, :halowidths)}}
_array(a::AbstractArray) = a
_array(a::NamedTuple) = a.A
_hw(a::AbstractArray) = Tuple(grid().halowidths)
_hw(a::NamedTuple) = Tuple(Int.(a.halowidths))
_devptr(A::CuArray) = reinterpret(Ptr{Cvoid}, pointer(A))
_devptr(A::Array) = Ptr{Cvoid}(pointer(A))          # must be page-locked (CUDA.pin / Mem.register); the library rejects pageable memory

# CellArray arguments (CellArrays.jl, loaded by the caller; recognised by their fields, so that this package does not
# depend on it): as the reference does (ref: src/shared.jl:246-250,287-289), a CellArray with B = 0 contributes one
# field per cell component (strided component views are copied by nobody: `field(A, i)` of a B = 0 CellArray is a dense
# N-dimensional array), one with B = 1 a single field whose elements are whole cells; other B are refused.
_is_cellarray(a) = hasproperty(a, :data) && hasproperty(a, :dims) && !(a isa AbstractArray{<:Number})
function _bitsarrays(A)
    T = eltype(A); B = typeof(A).parameters[3]
    if B == 0
        ncomp = T <: Number ? 1 : length(T)
        return [reshape(view(A.data, :, i, 1), size(A)) for i in 1:ncomp]
    elseif B == 1
        return [reshape(reinterpret(T, view(A.data, :)), size(A))]
    end
    error("only CellArrays with B=0 or B=1 are supported")
end
function _extract_all(A::Tuple)
    out = Any[]
    for a in A
        if a isa NamedTuple && _is_cellarray(a.A)
            append!(out, [(A=b, halowidths=a.halowidths) for b in _bitsarrays(a.A)])
        elseif _is_cellarray(a)
            append!(out, _bitsarrays(a))
        else
            push!(out, a)
        end
    end
    return Tuple(out)
end

"""
    update_halo!(A...; dims=(3,1,2))

Same call forms as ImplicitGlobalGrid.update_halo!: arrays and/or `(A=array, halowidths=(..))` named tuples.
One ccall; blocks until the halos are written.
"""
function update_halo!(A...; dims=(NDIMS, (1:NDIMS-1)...))
    gg = grid()
    A = _extract_all(A)
    arrays = map(_array, A)
    # type-level checks stay on the host side; the rest is checked by the library with the reference's messages
    for (i, a) in enumerate(arrays)
        (isbitstype(eltype(a)) || Base.isbitsunion(eltype(a))) || error("The field at position $i is not of bits type or 'is-bits' Union type.")
        (a isa CuArray || a isa Array) || error("The field at position $i does not have a supported array type.")
        typeof(a) == typeof(arrays[1]) || error("The field at position $i is of different type than the first field; make sure that in a same call all fields are of the same type.")
    end
    fields = [Field(_devptr(a), (size(a, 1), size(a, 2), size(a, 3)), _hw(f), Int32(sizeof(eltype(a))), Int32(ndims(a)))
              for (a, f) in zip(arrays, A)]
    d = Cint[dims...]
    GC.@preserve arrays begin
        check(ccall((:igg_update_halo, LIB), Cint, (Ptr{Cvoid}, Cint, Ptr{Field}, Ptr{Cint}, Cint, Ptr{Cvoid}, Cint),
                    gg.ctx, length(fields), fields, d, length(d), CUDA.stream().handle, 0))
    end
    return nothing
end

# ---- prepared calls and symmetric fields (library extensions, see include/igg_b200.h) ----------------------------
"`call = prepare_update_halo(A...; dims)`: validates once; `call()` repeats the update with one cheap ccall."
function prepare_update_halo(A...; dims=(NDIMS, (1:NDIMS-1)...), blocking::Bool=true)
    gg = grid()
    A = _extract_all(A)
    arrays = map(_array, A)
    fields = [Field(_devptr(a), (size(a, 1), size(a, 2), size(a, 3)), _hw(f), Int32(sizeof(eltype(a))), Int32(ndims(a)))
              for (a, f) in zip(arrays, A)]
    d = Cint[dims...]
    h = Ref{Ptr{Cvoid}}(C_NULL)
    check(ccall((:igg_prepare, LIB), Cint, (Ptr{Cvoid}, Cint, Ptr{Field}, Ptr{Cint}, Cint, Ref{Ptr{Cvoid}}),
                gg.ctx, length(fields), fields, d, length(d), h))
    handle = h[]
    return () -> (GC.@preserve arrays check(ccall((:igg_update_halo_prepared, LIB), Cint, (Ptr{Cvoid}, Ptr{Cvoid}, Cint),
                                                   handle, CUDA.stream().handle, blocking ? 0 : 1)); nothing)
end

"`alloc_field(T, dims...)`: a zeroed CuArray inside a symmetric allocation (collective): update_halo! may store into it from the neighbours directly."
function alloc_field(::Type{T}, dims::Integer...) where {T}
    gg = grid()
    p = Ref{Ptr{Cvoid}}(C_NULL)
    check(ccall((:igg_alloc_field, LIB), Cint, (Ptr{Cvoid}, Csize_t, Ref{Ptr{Cvoid}}), gg.ctx, prod(dims) * sizeof(T), p))
    return unsafe_wrap(CuArray, reinterpret(CuPtr{T}, p[]), dims; own=false)
end
free_field(A::CuArray) = (check(ccall((:igg_free_field, LIB), Cint, (Ptr{Cvoid}, Ptr{Cvoid}), grid().ctx, _devptr(A))); nothing)

# ---- gather! -----------------------------------------------------------------------------------------------------
"""
    gather!(A, A_global; root=0)

`A` / `A_global` may be `Array` or `CuArray`; `A_global` may be `nothing` on non-root ranks.
"""
function gather!(A::AbstractArray{T,N2}, A_global::Union{AbstractArray{T,N},Nothing}; root::Integer=0) where {T,N,N2}
    gg = grid()
    sA = Int64[size(A, 1), size(A, 2), size(A, 3)]
    if A_global === nothing
        check(ccall((:igg_gather, LIB), Cint, (Ptr{Cvoid}, Ptr{Cvoid}, Ptr{Int64}, Cint, Ptr{Cvoid}, Ptr{Int64}, Cint, Cint, Cint, Ptr{Cvoid}),
                    gg.ctx, _devptr(A), sA, N2, C_NULL, C_NULL, 0, sizeof(T), root, CUDA.stream().handle))
    else
        sG = Int64[size(A_global, 1), size(A_global, 2), size(A_global, 3)]
        GC.@preserve A A_global check(ccall((:igg_gather, LIB), Cint,
                    (Ptr{Cvoid}, Ptr{Cvoid}, Ptr{Int64}, Cint, Ptr{Cvoid}, Ptr{Int64}, Cint, Cint, Cint, Ptr{Cvoid}),
                    gg.ctx, _devptr(A), sA, N2, _devptr(A_global), sG, N, sizeof(T), root, CUDA.stream().handle))
    end
    return nothing
end

# ---- tools: global sizes, indices, coordinates (host-only index arithmetic) ---------------------------------------
_n_g(d::Int) = grid().nxyz_g[d]
function _n_g(d::Int, nA::Integer)
    gg = grid(); diff = nA - gg.nxyz[d]; p = gg.periods[d] > 0
    return gg.nxyz_g[d] + (p ? min(diff + gg.overlaps[d], 0) : diff)
end
nx_g() = _n_g(1); ny_g() = _n_g(2); nz_g() = _n_g(3)
nx_g(A::AbstractArray) = _n_g(1, size(A, 1)); ny_g(A::AbstractArray) = _n_g(2, size(A, 2)); nz_g(A::AbstractArray) = _n_g(3, size(A, 3))

function _i_g(d::Int, i::Integer, nA::Integer, wrap::Bool)
    gg = grid(); diff = nA - gg.nxyz[d]; olA = gg.overlaps[d] + diff; p = gg.periods[d] > 0
    ig = gg.coords[d] * (gg.nxyz[d] - gg.overlaps[d]) + i + (p ? -(olA ÷ 2) : 0)
    if p && wrap
        nga = _n_g(d, nA)
        ig > nga && (ig -= nga)
        ig < 1 && (ig += nga)
    end
    return ig
end
ix_g(i::Integer, A::AbstractArray; wrap_periodic::Bool=true) = _i_g(1, i, size(A, 1), wrap_periodic)
iy_g(i::Integer, A::AbstractArray; wrap_periodic::Bool=true) = _i_g(2, i, size(A, 2), wrap_periodic)
iz_g(i::Integer, A::AbstractArray; wrap_periodic::Bool=true) = _i_g(3, i, size(A, 3), wrap_periodic)

function _x_g(d::Int, i::Integer, dx::AbstractFloat, nA::Integer, wrap::Bool)
    gg = grid(); diff = nA - gg.nxyz[d]; p = gg.periods[d] > 0
    ng = gg.dims[d] * (gg.nxyz[d] - gg.overlaps[d]) + (p ? 0 : gg.overlaps[d])
    l = gg.origin_on_vertex ? dx * ng : dx * (ng - 1)
    x1 = gg.origin[d] + (gg.centerxyz[d] ? -l / 2 : 0.0) +
         dx * ((gg.origin_on_vertex ? 0.5 : 0.0) + (p ? 0.0 : -diff / 2) + ((p && isodd(diff)) ? -0.5 : 0.0))
    xs = x1:dx:(x1 + l + dx * diff)
    ig = _i_g(d, i, nA, wrap)
    ig > length(xs) && return xs[end] + dx * (ig - length(xs))
    ig < 1 && return xs[1] - dx * (1 - ig)
    return xs[ig]
end
x_g(i::Integer, dx::AbstractFloat, A::AbstractArray; wrap_periodic::Bool=true) = _x_g(1, i, dx, size(A, 1), wrap_periodic)
y_g(i::Integer, dy::AbstractFloat, A::AbstractArray; wrap_periodic::Bool=true) = _x_g(2, i, dy, size(A, 2), wrap_periodic)
z_g(i::Integer, dz::AbstractFloat, A::AbstractArray; wrap_periodic::Bool=true) = _x_g(3, i, dz, size(A, 3), wrap_periodic)

tic() = (CUDA.synchronize(); MPI.Barrier(grid().comm); T0[] = time(); nothing)
toc() = (CUDA.synchronize(); MPI.Barrier(grid().comm); time() - T0[])

end # module
