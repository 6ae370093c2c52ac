{-# LANGUAGE DataKinds #-}
{-# LANGUAGE LambdaCase #-}
{-# LANGUAGE RecordWildCards #-}
-- | Drop-in for @Graphics.Ray@ (Ray.hs:7-20): 'CameraSettings', 'defaultCameraSettings', 'ToRandom', 'raytrace',
-- 'readImage', 'writeImage', 'writeImageSqrt' keep their names, fields and types; 'raytrace' runs on a B200 through
-- librt_b200.so instead of the massiv worker pool.
--
-- Written against the reference's export list; NOT compiled in the authoring environment (no GHC).  The Python module
-- @raytrace_b200/ray.py@ is the same logic, line for line, and is what the test-suite runs.
module Graphics.Ray
  ( CameraSettings(..), defaultCameraSettings
  , ToRandom(toRandom), raytrace
  , readImage, writeImage, writeImageSqrt
  , module Graphics.Ray.Core
  , module Graphics.Ray.Geometry
  , module Graphics.Ray.Material
  , module Graphics.Ray.Texture
  , module Graphics.Ray.Noise
  ) where

import Graphics.Ray.Core
import Graphics.Ray.Geometry
import Graphics.Ray.Material
import Graphics.Ray.Texture
import Graphics.Ray.Noise
import Graphics.Ray.B200.FFI

import Linear (V2(V2), V3(V3), V4(V4), normalize)
import System.Random (StdGen, genWord64)
import Data.Massiv.Array (D, S, U, Ix2((:.)))
import qualified Data.Massiv.Array as A
import qualified Data.Massiv.Array.IO as I
import qualified Graphics.Pixel.ColorSpace as C
import Control.Monad.State (State)
import Control.Monad (forM, when, zipWithM_)
import Data.Functor.Identity (Identity, runIdentity)
import Foreign
import Foreign.C.Types (CFloat)
import System.IO.Unsafe (unsafePerformIO)

-- | Ray.hs:40-68, field for field.
data CameraSettings = CameraSettings
  { cs_center :: Point3
  , cs_lookAt :: Point3
  , cs_up :: Vec3
  , cs_vfov :: Double
  , cs_aspectRatio :: Double
  , cs_imageWidth :: Int
  , cs_samplesPerPixel :: Int
  , cs_maxRecursionDepth :: Int
  , cs_background :: Ray -> Color
  , cs_defocusAngle :: Double
  , cs_focusDist :: Double
  , cs_redirectTargets :: [(Double, Point3, Vec3, Vec3)]
  }

-- | Ray.hs:84-98.
defaultCameraSettings :: CameraSettings
defaultCameraSettings = CameraSettings
  { cs_center = V3 0 0 0, cs_lookAt = V3 0 0 (-1), cs_up = V3 0 1 0, cs_vfov = pi / 2, cs_aspectRatio = 1
  , cs_imageWidth = 100, cs_samplesPerPixel = 10, cs_maxRecursionDepth = 10, cs_background = const (V3 1 1 1)
  , cs_defocusAngle = 0, cs_focusDist = 10, cs_redirectTargets = [] }

-- | Kept so that both @Geometry Identity Material@ and @Geometry (State StdGen) Material@ (volumes) are accepted
-- (Ray.hs:100-109); the backend does not need the method.
class ToRandom m where
  toRandom :: m a -> State StdGen a
instance ToRandom Identity where
  toRandom = pure . runIdentity
instance ToRandom (State StdGen) where
  toRandom = id

t3 :: V3 Double -> (Double, Double, Double)
t3 (V3 x y z) = (x, y, z)

-- | @cs_background@ keeps its closure type; it is classified by probing: constant, or affine in @dir.y@ (the two
-- backgrounds of the reference's test program, test/Main.hs:19-28).  Anything else is a user-defined closure.
classifyBackground :: (Ray -> Color) -> Either String (Int32, Color, Color)
classifyBackground bg
  | all (close up) vals && close up dn = Right (1, up, up)
  | and [ close (lerp d) v | (d, v) <- zip dirs vals ] = Right (2, dn, up)
  | otherwise = Left "cs_background is neither constant nor affine in the ray direction's y component; user-defined backgrounds are rejected (no CPU fallback)"
  where
    up = bg (Ray (V3 0 0 0) (V3 0 1 0))
    dn = bg (Ray (V3 0 0 0) (V3 0 (-1) 0))
    dirs = [ normalize (V3 (sin (1.7 * k)) (cos (2.3 * k + 0.4)) (sin (0.9 * k + 1.1))) | k <- [1 .. 64] ]
    origins = [ V3 (37 * sin k) (91 * cos k) (53 * sin (2 * k)) | k <- [1 .. 64] ]
    vals = [ bg (Ray o d) | (o, d) <- zip origins dirs ]
    lerp (V3 _ y _) = let a = 0.5 * (y + 1) in fmap (* (1 - a)) dn + fmap (* a) up
    close (V3 a b c) (V3 x y z) = maximum [abs (a - x), abs (b - y), abs (c - z)] <= 1e-9

-- Post-order serialisation of the description tree into the scene handle: one function per payload type
-- (Material at and above `<$`, () below it).
nodeM :: Ptr RtScene -> (Material -> IO Word32) -> GeometryRep Material -> IO Word32
nodeM scene matId = \case
  WithPayload m g -> do
    c <- nodeU scene g
    mid <- matId m
    emitNode scene 12 [c] (\p -> pokeNodeMaterial p mid)
  Group gs -> mapM (nodeM scene matId) gs >>= \cs -> emitNode scene 6 cs (const (pure ()))
  BvhNode l r -> mapM (nodeM scene matId) [l, r] >>= \cs -> emitNode scene 7 cs (const (pure ()))
  BvhTree gs -> mapM (nodeM scene matId) gs >>= \cs -> emitNode scene 8 cs (const (pure ()))
  Transform m g -> nodeM scene matId g >>= \c -> emitNode scene 9 [c] (\p -> pokeNodeM p (flatten44 m))
  Moving a b g -> nodeM scene matId g >>= \c -> emitNode scene 10 [c] (\p -> pokeNodeV p 0 (t3 a) >> pokeNodeV p 1 (t3 b))
  _ -> ioError (userError "raytrace: world is a Geometry m () — attach materials with (<$)")

nodeU :: Ptr RtScene -> GeometryRep () -> IO Word32
nodeU scene = \case
  Sphere c r -> leaf 1 (\p -> pokeNodeV p 0 (t3 c) >> pokeNodeS p 0 r)
  Parallelogram q u v -> leaf 2 (\p -> zipWithM_ (pokeNodeV p) [0 ..] (map t3 [q, u, v]))
  Triangle a b c -> leaf 3 (\p -> do
    zipWithM_ (pokeNodeV p) [0 ..] (map (t3 . fst) [a, b, c])
    zipWithM_ (pokeNodeUV p) [0 ..] [ (u, v) | (_, V2 u v) <- [a, b, c] ])
  Cuboid ((x0, x1), (y0, y1), (z0, z1)) -> leaf 4 (\p -> pokeNodeV p 0 (x0, y0, z0) >> pokeNodeV p 1 (x1, y1, z1))
  TriMesh (Mesh vs ts fs) ->
    withArrayLen (concat [ [x, y, z] | V3 x y z <- A.toList vs ]) $ \nv pv ->
    withArrayLen (concat [ [u, v] | V2 u v <- A.toList ts ]) $ \nt pt ->
    withArrayLen (concat [ [i, ix a, j, ix b, k, ix c] | V3 (i', a) (j', b) (k', c) <- fs
                                                         , let (i, j, k) = (fromIntegral i', fromIntegral j', fromIntegral k') ]) $ \nf pf ->
      leaf 5 (\p -> do
        pokeNodeNVerts p (fromIntegral (nv `div` 3)) >> pokeNodeNUvs p (fromIntegral (nt `div` 2)) >> pokeNodeNFaces p (fromIntegral (nf `div` 6))
        pokeNodeVerts p pv >> pokeNodeUvs p pt >> pokeNodeFaces p pf)
  Group gs -> mapM (nodeU scene) gs >>= \cs -> emitNode scene 6 cs (const (pure ()))
  BvhNode l r -> mapM (nodeU scene) [l, r] >>= \cs -> emitNode scene 7 cs (const (pure ()))
  BvhTree gs -> mapM (nodeU scene) gs >>= \cs -> emitNode scene 8 cs (const (pure ()))
  Transform m g -> nodeU scene g >>= \c -> emitNode scene 9 [c] (\p -> pokeNodeM p (flatten44 m))
  Moving a b g -> nodeU scene g >>= \c -> emitNode scene 10 [c] (\p -> pokeNodeV p 0 (t3 a) >> pokeNodeV p 1 (t3 b))
  ConstantMedium d b -> nodeU scene b >>= \c -> emitNode scene 11 [c] (\p -> pokeNodeS p 0 d)
  WithPayload () g -> nodeU scene g
  UserDefined _ -> leaf 14 (const (pure ()))           -- RT_NODE_USER → RT_ERR_USER_DEFINED at commit
  where leaf k = emitNode scene k []
        ix = maybe (-1) fromIntegral

emitNode :: Ptr RtScene -> Int32 -> [Word32] -> (Ptr () -> IO ()) -> IO Word32
emitNode scene kind children fill = allocaBytes nodeSize $ \p -> withArrayLen children $ \n pc -> do
  fillBytes p 0 nodeSize
  pokeNodeKind p kind
  pokeNodeNChildren p (fromIntegral n)
  pokeNodeChildren p pc
  fill p
  alloca $ \o -> c_add_node scene p o >>= check >> peek o

flatten44 :: V4 (V4 Double) -> [Double]
flatten44 (V4 a b c d) = concatMap (\(V4 x y z w) -> [x, y, z, w]) [a, b, c, d]

-- | @raytrace settings world gen@ (Ray.hs:121): a pure function of its arguments, as in the reference.  The 64-bit
-- Philox key is drawn from @gen@; the sample stream is NOT the reference's SplitMix stream (parity is statistical).
-- A constantMedium under the payload-less part of the tree is also a 'WithPayload' at its root, so one entry suffices.
raytrace :: ToRandom m => CameraSettings -> Geometry m Material -> StdGen -> A.Matrix D Color
raytrace CameraSettings{..} world gen = unsafePerformIO $ do
  (bgKind, bg0, bg1) <- either (ioError . userError) pure (classifyBackground cs_background)
  scene <- alloca $ \pp -> c_scene_create pp >>= check >> peek pp
  let matId = serialiseMaterial scene
  root <- nodeM scene matId (geometryRep world)
  -- the package's own Perlin tables (Noise.hs:52-92) pin noise parity: pass them instead of the library defaults
  withArray (map fromIntegral noisePermX) $ \px -> withArray (map fromIntegral noisePermY) $ \py ->
    withArray (map fromIntegral noisePermZ) $ \pz -> withArray (concat [ [x, y, z] | V3 x y z <- noiseGradients ]) $ \pg ->
      c_set_noise scene px py pz pg >>= check
  c_scene_commit scene root >>= check
  let w = cs_imageWidth
      h = round (fromIntegral w / cs_aspectRatio) :: Int          -- round-half-even (Ray.hs:123)
      targets = [ RedirectTarget p (t3 q) (t3 a) (t3 b) | (p, q, a, b) <- cs_redirectTargets ]
      (seed, _) = genWord64 gen
  -- page-locked result buffer from the library's recycling pool: rt_render's copy engine writes it directly
  -- (a pageable mallocForeignPtrArray works too, through one staging memcpy)
  buf <- alloca $ \pp -> do
    c_host_alloc (fromIntegral (w * h * 3 * sizeOf (0 :: CFloat))) pp >>= check
    peek pp >>= newForeignPtr p_host_free . castPtr :: IO (ForeignPtr CFloat)
  withArrayLen targets $ \nt pt -> allocaBytes cameraSize $ \cam -> withForeignPtr buf $ \o -> do
    pokeCamera cam Camera
      { camCenter = t3 cs_center, camLookAt = t3 cs_lookAt, camUp = t3 cs_up, camVfov = cs_vfov, camAspect = cs_aspectRatio
      , camDefocus = cs_defocusAngle, camFocus = cs_focusDist, camBg0 = t3 bg0, camBg1 = t3 bg1, camBgKind = bgKind
      , camWidth = fromIntegral w, camSpp = fromIntegral cs_samplesPerPixel, camDepth = fromIntegral cs_maxRecursionDepth
      , camTargets = pt, camNTargets = fromIntegral nt }
    -- every visible GPU renders its share of the samples and DMA-copies its slice of the reduced image into the buffer
    -- (one GPU: the same as rt_render scene 0 ...)
    c_render_multi allGpus scene cam seed o nullPtr nullPtr >>= check
  _ <- c_scene_destroy scene
  px <- withForeignPtr buf (peekArray (w * h * 3))
  let img = A.fromList A.Seq (map realToFrac px) :: A.Vector U Double
  pure $ A.makeArray A.Par (A.Sz2 h w) $ \(j :. i) ->           -- index (j :. i), row 0 = top (Ray.hs:238)
    let k = (j * w + i) * 3 in V3 (img A.! k) (img A.! (k + 1)) (img A.! (k + 2))

-- | The process-wide context over all sm_100 GPUs (rt_context_create NULL 0), made on first use and kept until exit.
allGpus :: Ptr RtContext
allGpus = unsafePerformIO $ alloca $ \pp -> c_context_create nullPtr 0 pp >>= check >> peek pp
{-# NOINLINE allGpus #-}

serialiseMaterial :: Ptr RtScene -> Material -> IO Word32
serialiseMaterial scene m = do
  let tex t = case textureRep t of
        TexConstant c -> addT (base { tdKind = 1, tdC0 = t3 c })
        TexChecker nu nv c0 c1 -> addT (base { tdKind = 2, tdI0 = fromIntegral nu, tdI1 = fromIntegral nv, tdC0 = t3 c0, tdC1 = t3 c1 })
        TexImage img -> do
          let A.Sz2 hh ww = A.size img
          withArray (concat [ map realToFrac [x, y, z] | V3 x y z <- A.toList img ] :: [CFloat]) $ \p ->
            addT (base { tdKind = 3, tdWidth = fromIntegral ww, tdHeight = fromIntegral hh, tdImage = p })
        TexNoise k f s c0 c1 -> addT (base { tdKind = 4, tdI0 = fromIntegral k, tdFreq = f, tdShift = t3 s, tdC0 = t3 c0, tdC1 = t3 c1 })
        TexMarble d f s -> addT (base { tdKind = 5, tdFreq = f, tdShift = t3 s, tdDir = t3 d })
        TexUser _ -> addT (base { tdKind = 6 })
      base = TextureDesc 0 0 0 0 0 0 (0, 0, 0) (0, 0, 0) (0, 0, 0) (0, 0, 0) nullPtr
      addT d = with d $ \p -> alloca $ \o -> c_add_texture scene p o >>= check >> peek o
      mat k mt x = do
        tid <- maybe (pure 0) tex mt
        with (MaterialDesc k tid x) $ \p -> alloca $ \o -> c_add_material scene p o >>= check >> peek o
  case m of
    Material _ -> mat 11 Nothing 0
    LightSource t -> mat 1 (Just t) 0
    PitchBlack -> mat 2 Nothing 0
    Lambertian t -> mat 3 (Just t) 0
    LommelSeeliger t -> mat 4 (Just t) 0
    Mirror t -> mat 5 (Just t) 0
    Metal f t -> mat 6 (Just t) f
    Dielectric i -> mat 7 Nothing i
    Transparent t -> mat 8 (Just t) 0
    Isotropic t -> mat 9 (Just t) 0
    Anisotropic g t -> mat 10 (Just t) g

-- | Ray.hs:241-245: decode, sRGB → linear.
readImage :: FilePath -> IO (A.Matrix U Color)
readImage path = do
  img <- I.readImageAuto path :: IO (A.Matrix S (C.Pixel (C.SRGB 'C.Linear) Double))
  pure $ A.compute $ A.map (\(C.PixelSRGB r g b) -> V3 r g b) img

-- | Ray.hs:248-252 / 256-260: the OETF + 8-bit quantisation run on the device (rt_tonemap_u8: mode 0 = sRGB, 1 = sqrt).
writeImage, writeImageSqrt :: A.Source r Color => FilePath -> A.Matrix r Color -> IO ()
writeImage = writeWith 0
writeImageSqrt = writeWith 1

writeWith :: A.Source r Color => Int32 -> FilePath -> A.Matrix r Color -> IO ()
writeWith mode path m = do
  let A.Sz2 h w = A.size m
      n = h * w
  bytes <- withArray (concat [ map realToFrac [x, y, z] | V3 x y z <- A.toList m ] :: [CFloat]) $ \pin ->
    allocaArray (3 * n) $ \pout -> do
      c_tonemap_u8 0 pin (fromIntegral n) mode pout >>= check
      peekArray (3 * n) pout
  let px = A.fromLists' A.Seq [ [ C.PixelSRGB (bytes !! k) (bytes !! (k + 1)) (bytes !! (k + 2)) | i <- [0 .. w - 1], let k = 3 * (j * w + i) ]
                              | j <- [0 .. h - 1] ] :: A.Matrix S (C.Pixel (C.SRGB 'C.NonLinear) Word8)
  I.writeImageAuto path px
