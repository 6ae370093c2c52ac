{-# LANGUAGE ForeignFunctionInterface #-}
-- | Raw binding of @include/rt_b200.h@ (librt_b200.so).  One @foreign import@ per entry point; every call that can
-- touch the GPU is @safe@ (a render runs for milliseconds to minutes and must not block the other Haskell threads;
-- the library sets the CUDA device explicitly in every call, so migrating OS threads is fine).
--
-- NOT COMPILED IN THE AUTHORING ENVIRONMENT (no GHC there); the same ABI is exercised by the ctypes binding
-- @raytrace_b200/_ffi.py@, whose struct layouts this file mirrors field for field.
module Graphics.Ray.B200.FFI where

import Foreign
import Foreign.C.String (CString, peekCString)
import Foreign.C.Types

#include "rt_b200.h"

data RtScene   -- opaque handle
data RtMesh    -- opaque handle
data RtContext -- opaque handle: a set of GPUs driven by one process

-- | Status codes (@RtStatus@).
rtOk, rtErrEmptyBvh, rtErrUserDefined :: Int32
rtOk             = #{const RT_OK}
rtErrEmptyBvh    = #{const RT_ERR_EMPTY_BVH}
rtErrUserDefined = #{const RT_ERR_USER_DEFINED}

-- ---------------------------------------------------------------------------------------------------------------
-- POD structs
-- ---------------------------------------------------------------------------------------------------------------
data TextureDesc = TextureDesc
  { tdKind :: Int32, tdI0, tdI1 :: Int32, tdWidth, tdHeight :: Int32, tdFreq :: Double
  , tdC0, tdC1, tdShift, tdDir :: (Double, Double, Double), tdImage :: Ptr CFloat }

pokeD3 :: Ptr a -> Int -> (Double, Double, Double) -> IO ()
pokeD3 p off (x, y, z) = pokeByteOff p off x >> pokeByteOff p (off + 8) y >> pokeByteOff p (off + 16) z

instance Storable TextureDesc where
  sizeOf _ = #{size RtTextureDesc}
  alignment _ = #{alignment RtTextureDesc}
  peek _ = error "RtTextureDesc is write-only"
  poke p t = do
    fillBytes p 0 #{size RtTextureDesc}
    #{poke RtTextureDesc, kind} p (tdKind t)
    #{poke RtTextureDesc, i0} p (tdI0 t)
    #{poke RtTextureDesc, i1} p (tdI1 t)
    #{poke RtTextureDesc, width} p (tdWidth t)
    #{poke RtTextureDesc, height} p (tdHeight t)
    #{poke RtTextureDesc, freq} p (tdFreq t)
    pokeD3 p #{offset RtTextureDesc, c0} (tdC0 t)
    pokeD3 p #{offset RtTextureDesc, c1} (tdC1 t)
    pokeD3 p #{offset RtTextureDesc, shift} (tdShift t)
    pokeD3 p #{offset RtTextureDesc, dir} (tdDir t)
    #{poke RtTextureDesc, image} p (tdImage t)

data MaterialDesc = MaterialDesc { mdKind :: Int32, mdTexture :: Word32, mdParam :: Double }

instance Storable MaterialDesc where
  sizeOf _ = #{size RtMaterialDesc}
  alignment _ = #{alignment RtMaterialDesc}
  peek _ = error "RtMaterialDesc is write-only"
  poke p m = do
    #{poke RtMaterialDesc, kind} p (mdKind m)
    #{poke RtMaterialDesc, texture} p (mdTexture m)
    #{poke RtMaterialDesc, param} p (mdParam m)

-- | @RtNodeDesc@ is filled field by field by 'Graphics.Ray.Geometry.serialise' through these offsets.
nodeSize :: Int
nodeSize = #{size RtNodeDesc}

pokeNodeKind :: Ptr a -> Int32 -> IO ()
pokeNodeKind = #{poke RtNodeDesc, kind}
pokeNodeMaterial, pokeNodeNChildren, pokeNodeNVerts, pokeNodeNUvs, pokeNodeNFaces :: Ptr a -> Word32 -> IO ()
pokeNodeMaterial  = #{poke RtNodeDesc, material}
pokeNodeNChildren = #{poke RtNodeDesc, n_children}
pokeNodeNVerts    = #{poke RtNodeDesc, n_verts}
pokeNodeNUvs      = #{poke RtNodeDesc, n_uvs}
pokeNodeNFaces    = #{poke RtNodeDesc, n_faces}
pokeNodeChildren :: Ptr a -> Ptr Word32 -> IO ()
pokeNodeChildren = #{poke RtNodeDesc, children}
pokeNodeVerts, pokeNodeUvs :: Ptr a -> Ptr Double -> IO ()
pokeNodeVerts = #{poke RtNodeDesc, verts}
pokeNodeUvs   = #{poke RtNodeDesc, uvs}
pokeNodeFaces :: Ptr a -> Ptr Int32 -> IO ()
pokeNodeFaces = #{poke RtNodeDesc, faces}
-- | vector slot k (0..2), scalar slot k (0..1), uv slot k (0..2), matrix (16 doubles, row-major)
pokeNodeV :: Ptr a -> Int -> (Double, Double, Double) -> IO ()
pokeNodeV p k = pokeD3 p (#{offset RtNodeDesc, v} + 24 * k)
pokeNodeS :: Ptr a -> Int -> Double -> IO ()
pokeNodeS p k = pokeByteOff p (#{offset RtNodeDesc, s} + 8 * k)
pokeNodeUV :: Ptr a -> Int -> (Double, Double) -> IO ()
pokeNodeUV p k (u, v) = pokeByteOff p (#{offset RtNodeDesc, uv} + 16 * k) u >> pokeByteOff p (#{offset RtNodeDesc, uv} + 16 * k + 8) v
pokeNodeM :: Ptr a -> [Double] -> IO ()
pokeNodeM p = pokeArray (p `plusPtr` #{offset RtNodeDesc, m})

data RedirectTarget = RedirectTarget Double (Double, Double, Double) (Double, Double, Double) (Double, Double, Double)

instance Storable RedirectTarget where
  sizeOf _ = #{size RtRedirectTarget}
  alignment _ = #{alignment RtRedirectTarget}
  peek _ = error "RtRedirectTarget is write-only"
  poke p (RedirectTarget pr q s0 s1) = do
    #{poke RtRedirectTarget, p} p pr
    pokeD3 p #{offset RtRedirectTarget, q} q
    pokeD3 p #{offset RtRedirectTarget, s0} s0
    pokeD3 p #{offset RtRedirectTarget, s1} s1

-- | @RtCamera@: CameraSettings field for field, background already classified.
data Camera = Camera
  { camCenter, camLookAt, camUp :: (Double, Double, Double)
  , camVfov, camAspect, camDefocus, camFocus :: Double
  , camBg0, camBg1 :: (Double, Double, Double), camBgKind :: Int32
  , camWidth, camSpp, camDepth :: Int32
  , camTargets :: Ptr RedirectTarget, camNTargets :: Int32 }

cameraSize :: Int
cameraSize = #{size RtCamera}

pokeCamera :: Ptr a -> Camera -> IO ()
pokeCamera p c = do
  fillBytes p 0 cameraSize
  pokeD3 p #{offset RtCamera, center} (camCenter c)
  pokeD3 p #{offset RtCamera, look_at} (camLookAt c)
  pokeD3 p #{offset RtCamera, up} (camUp c)
  #{poke RtCamera, vfov} p (camVfov c)
  #{poke RtCamera, aspect_ratio} p (camAspect c)
  #{poke RtCamera, defocus_angle} p (camDefocus c)
  #{poke RtCamera, focus_dist} p (camFocus c)
  pokeD3 p #{offset RtCamera, bg_c0} (camBg0 c)
  pokeD3 p #{offset RtCamera, bg_c1} (camBg1 c)
  #{poke RtCamera, bg_kind} p (camBgKind c)
  #{poke RtCamera, image_width} p (camWidth c)
  #{poke RtCamera, samples_per_pixel} p (camSpp c)
  #{poke RtCamera, max_recursion_depth} p (camDepth c)
  #{poke RtCamera, n_redirect_targets} p (camNTargets c)
  #{poke RtCamera, redirect_targets} p (camTargets c)

-- ---------------------------------------------------------------------------------------------------------------
-- entry points
-- ---------------------------------------------------------------------------------------------------------------
foreign import ccall unsafe "rt_version"                c_rt_version     :: IO Word32
foreign import ccall unsafe "rt_last_error"             c_rt_last_error  :: IO CString
foreign import ccall safe   "rt_device_count"           c_device_count   :: Ptr Int32 -> IO Int32
foreign import ccall safe   "rt_scene_create"           c_scene_create   :: Ptr (Ptr RtScene) -> IO Int32
foreign import ccall safe   "rt_scene_destroy"          c_scene_destroy  :: Ptr RtScene -> IO Int32
foreign import ccall safe   "rt_scene_add_texture"      c_add_texture    :: Ptr RtScene -> Ptr TextureDesc -> Ptr Word32 -> IO Int32
foreign import ccall safe   "rt_scene_add_material"     c_add_material   :: Ptr RtScene -> Ptr MaterialDesc -> Ptr Word32 -> IO Int32
foreign import ccall safe   "rt_scene_add_node"         c_add_node       :: Ptr RtScene -> Ptr () -> Ptr Word32 -> IO Int32
foreign import ccall safe   "rt_scene_set_noise_tables" c_set_noise      :: Ptr RtScene -> Ptr Int32 -> Ptr Int32 -> Ptr Int32 -> Ptr Double -> IO Int32
foreign import ccall safe   "rt_scene_commit"           c_scene_commit   :: Ptr RtScene -> Word32 -> IO Int32
foreign import ccall safe   "rt_scene_upload"           c_scene_upload   :: Ptr RtScene -> Int32 -> IO Int32
foreign import ccall safe   "rt_render"                 c_render         :: Ptr RtScene -> Int32 -> Ptr () -> Word64 -> Word64 -> Word64
                                                                         -> Ptr CFloat -> Ptr CFloat -> Ptr () -> IO Int32
-- one call, the whole machine (Ray.hs:121,238): a context over every visible GPU; rt_render_multi splits the samples per pixel
-- across its devices and assembles the image inside the library (reduce-scatter over NVLink, every GPU copies its slice)
foreign import ccall safe   "rt_context_create"         c_context_create :: Ptr Int32 -> Int32 -> Ptr (Ptr RtContext) -> IO Int32   -- nullPtr 0 = all sm_100 GPUs
foreign import ccall safe   "rt_context_destroy"        c_context_destroy :: Ptr RtContext -> IO Int32
foreign import ccall safe   "rt_render_multi"           c_render_multi   :: Ptr RtContext -> Ptr RtScene -> Ptr () -> Word64
                                                                         -> Ptr CFloat -> Ptr CFloat -> Ptr () -> IO Int32
-- a whole marshalled world in one call: three arrays of records, ids consecutive from the scene's current counts (a world is
-- immutable, so its arrays can be marshalled once and handed to every scene)
foreign import ccall safe   "rt_scene_add_batch"        c_scene_add_batch :: Ptr RtScene -> Ptr () -> Word32 -> Ptr () -> Word32 -> Ptr () -> Word32
                                                                         -> Ptr Word32 -> IO Int32
foreign import ccall safe   "rt_tonemap_u8_static"      c_tonemap_u8_static :: Int32 -> Ptr CFloat -> Int32 -> Int32 -> Ptr Word8 -> Ptr () -> IO Int32
-- tuning / A-B options are a struct (RtOptions: gpu_build, material_sampling, ...), never environment variables
foreign import ccall unsafe "rt_options_init"           c_options_init   :: Ptr () -> IO Int32
foreign import ccall safe   "rt_scene_set_options"      c_scene_set_options :: Ptr RtScene -> Ptr () -> IO Int32
foreign import ccall safe   "rt_host_alloc"             c_host_alloc     :: CSize -> Ptr (Ptr ()) -> IO Int32
foreign import ccall safe   "&rt_host_free"             p_host_free      :: FunPtr (Ptr CFloat -> IO ())   -- ForeignPtr finaliser (status ignored)
foreign import ccall safe   "rt_probe_primary"          c_probe_primary  :: Ptr RtScene -> Int32 -> Ptr () -> Double -> Ptr Word32 -> Ptr CFloat -> IO Int32
foreign import ccall safe   "rt_probe_surface"          c_probe_surface  :: Ptr RtScene -> Int32 -> Ptr () -> Double -> Ptr CFloat -> IO Int32
foreign import ccall safe   "rt_tonemap_u8"             c_tonemap_u8     :: Int32 -> Ptr CFloat -> Int32 -> Int32 -> Ptr Word8 -> IO Int32
foreign import ccall safe   "rt_obj_parse"              c_obj_parse      :: CString -> CSize -> Ptr (Ptr RtMesh) -> IO Int32
foreign import ccall unsafe "rt_mesh_sizes"             c_mesh_sizes     :: Ptr RtMesh -> Ptr Word32 -> Ptr Word32 -> Ptr Word32 -> IO Int32
foreign import ccall unsafe "rt_mesh_data"              c_mesh_data      :: Ptr RtMesh -> Ptr (Ptr Double) -> Ptr (Ptr Double) -> Ptr (Ptr Int32) -> IO Int32
foreign import ccall unsafe "rt_mesh_destroy"           c_mesh_destroy   :: Ptr RtMesh -> IO Int32

-- | Turn a non-zero status into an 'IOError' carrying the library's message.  @bvhTree []@ keeps the reference's text
-- (@"bvhTree: empty list"@, Geometry.hs:371); user-defined closures say so explicitly (there is no CPU fallback).
check :: Int32 -> IO ()
check 0 = pure ()
check _ = c_rt_last_error >>= peekCString >>= ioError . userError
