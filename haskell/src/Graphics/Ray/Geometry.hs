{-# LANGUAGE LambdaCase #-}
-- | Drop-in for @Graphics.Ray.Geometry@ (Geometry.hs:5-16): the same combinators with the same types, but a
-- @Geometry m a@ is a DESCRIPTION TREE (one constructor per combinator, mirroring @RtNodeKind@ of rt_b200.h) instead of
-- a closure, so that 'Graphics.Ray.raytrace' can hand it to the CUDA backend.  The raw 'Geometry' constructor and
-- 'planeShape' (arbitrary predicate) are kept for source compatibility and rejected at render time with an explicit
-- error — there is no CPU fallback.
--
-- Written against the reference's export list; NOT compiled in the authoring environment (no GHC).
module Graphics.Ray.Geometry
  ( Geometry(Geometry), boundingBox, pureGeometry, transform, moving
  , sphere, planeShape, parallelogram, cuboid, triangle, triangleMesh, constantMedium
  , Mesh(Mesh), transformVertices, parseObj, readObj
  , group, bvhNode, bvhTree
  , translate, rotateX, rotateY, rotateZ, scale
  , GeometryRep(..), geometryRep
  ) where

import Graphics.Ray.Core
import Graphics.Ray.B200.FFI

import Linear (V2(V2), V3(V3), V4(V4), M44, (!*), (^+^), (^-^))
import qualified Data.Massiv.Array as A
import Data.Massiv.Array (U)
import Control.Monad.State (State)
import System.Random (StdGen)
import Data.Functor.Identity (Identity)
import Foreign
import Foreign.C.String (withCStringLen, peekCString)
import System.IO.Unsafe (unsafePerformIO)

-- | Polygonal mesh: vertex positions, texture coordinates, triangles indexing into both (Geometry.hs:179-184).
data Mesh = Mesh (A.Vector U Point3) (A.Vector U (V2 Double)) [V3 (Int, Maybe Int)]

-- | What the backend sees.  The type parameters are phantom apart from the payload of 'WithPayload'.
data GeometryRep a
  = Sphere Point3 Double
  | Parallelogram Point3 Vec3 Vec3
  | Triangle (Point3, V2 Double) (Point3, V2 Double) (Point3, V2 Double)
  | Cuboid Box
  | TriMesh Mesh
  | Group [GeometryRep a]
  | BvhNode (GeometryRep a) (GeometryRep a)
  | BvhTree [GeometryRep a]
  | Transform (M44 Double) (GeometryRep a)
  | Moving Vec3 Vec3 (GeometryRep a)
  | ConstantMedium Double (GeometryRep ())       -- ^ boundary is always pure geometry (Geometry.hs:301)
  | WithPayload a (GeometryRep ())               -- ^ @x <$ g@: the OUTERMOST payload wins, as with fmap (const x)
  | UserDefined Box                              -- ^ closure-built geometry: rejected at commit time

-- | Same name as the reference's constructor; building geometry from a raw closure yields 'UserDefined'.
data Geometry m a
  = Geometry Box (Double -> Ray -> Interval -> m (Maybe (HitRecord, a)))
  | Reified (GeometryRep a)

geometryRep :: Geometry m a -> GeometryRep a
geometryRep (Reified r) = r
geometryRep (Geometry box _) = UserDefined box

-- | Only @x <$ g@ (constant payload) can be expressed as data; it is the only use the library and its users make of the
-- Functor instance (Geometry.hs:44-47).  A general 'fmap' needs the old payload, which exists only on the GPU.
instance Functor (Geometry m) where
  fmap _ _ = error "Graphics.Ray.Geometry (B200 backend): only (<$) is supported on Geometry"
  x <$ g = Reified (WithPayload x (strip (geometryRep g)))
    where
      strip :: GeometryRep b -> GeometryRep ()
      strip = \case
        Sphere c r -> Sphere c r
        Parallelogram q u v -> Parallelogram q u v
        Triangle a b c -> Triangle a b c
        Cuboid b -> Cuboid b
        TriMesh m -> TriMesh m
        Group gs -> Group (map strip gs)
        BvhNode l r -> BvhNode (strip l) (strip r)
        BvhTree gs -> BvhTree (map strip gs)
        Transform m g' -> Transform m (strip g')
        Moving a b g' -> Moving a b (strip g')
        ConstantMedium d b -> ConstantMedium d b
        WithPayload _ g' -> g'
        UserDefined b -> UserDefined b

pureGeometry :: Applicative m => Geometry Identity a -> Geometry m a
pureGeometry = Reified . geometryRep

-- | Bounding boxes follow the reference's formulas exactly (they decide the 'bvhTree' split, Geometry.hs:369-377); the
-- backend recomputes the same boxes from the description tree, this function exists for user code.
boundingBox :: Geometry m a -> Box
boundingBox = go . geometryRep
  where
    go :: GeometryRep b -> Box
    go = \case
      Sphere c r -> fromCorners (c ^-^ V3 r r r) (c ^+^ V3 r r r)                          -- Geometry.hs:60-61
      Parallelogram q u v -> padBox 0.0001 (boxHull [q, q ^+^ u, q ^+^ v, q ^+^ u ^+^ v])  -- Geometry.hs:144,151
      Triangle (a, _) (b, _) (c, _) -> padBox 0.0001 (boxHull [a, b, c])
      Cuboid b -> padBox 0.0001 b
      TriMesh (Mesh vs _ fs) -> boxJoin [ padBox 0.0001 (boxHull [vs A.! i, vs A.! j, vs A.! k]) | V3 (i, _) (j, _) (k, _) <- fs ]
      Group gs -> boxJoin (map go gs)
      BvhNode l r -> boxJoin [go l, go r]
      BvhTree gs -> boxJoin (map go gs)
      Transform m g -> boxHull [ dropW (m !* V4 x y z 1) | V3 x y z <- allCorners (go g) ] -- Geometry.hs:386-387
      Moving a b g -> boxJoin [shiftBox a (go g), shiftBox b (go g)]                        -- Geometry.hs:455
      ConstantMedium _ b -> go b
      WithPayload _ g -> go g
      UserDefined b -> b
    dropW (V4 x y z _) = V3 x y z

sphere :: Point3 -> Double -> Geometry Identity ()
sphere c r = Reified (Sphere c r)

-- | Arbitrary predicate / uv closures cannot cross the C ABI: the result is a 'UserDefined' node.  'parallelogram' and
-- 'triangle', the library's own uses of planeShape, are constructors of their own.
planeShape :: Point3 -> Vec3 -> Vec3 -> (Double -> Double -> Bool) -> (Double -> Double -> V2 Double) -> Box -> Geometry Identity ()
planeShape _ _ _ _ _ box = Reified (UserDefined (padBox 0.0001 box))

parallelogram :: Point3 -> Vec3 -> Vec3 -> Geometry Identity ()
parallelogram q u v = Reified (Parallelogram q u v)

cuboid :: Box -> Geometry Identity ()
cuboid = Reified . Cuboid

triangle :: (Point3, V2 Double) -> (Point3, V2 Double) -> (Point3, V2 Double) -> Geometry Identity ()
triangle a b c = Reified (Triangle a b c)

triangleMesh :: Mesh -> Geometry Identity ()
triangleMesh (Mesh _ _ []) = error "bvhTree: empty list"      -- Geometry.hs:371 through :294
triangleMesh m = Reified (TriMesh m)

constantMedium :: Double -> Geometry Identity () -> Geometry (State StdGen) ()
constantMedium density boundary = Reified (ConstantMedium density (geometryRep boundary))

group :: Monad m => [Geometry m a] -> Geometry m a
group = Reified . Group . map geometryRep

bvhNode :: Monad m => Geometry m a -> Geometry m a -> Geometry m a
bvhNode l r = Reified (BvhNode (geometryRep l) (geometryRep r))

bvhTree :: Monad m => [Geometry m a] -> Geometry m a
bvhTree [] = error "bvhTree: empty list"                       -- Geometry.hs:371
bvhTree gs = Reified (BvhTree (map geometryRep gs))

transform :: Functor m => M44 Double -> Geometry m a -> Geometry m a
transform m = Reified . Transform m . geometryRep

moving :: Functor m => Vec3 -> Vec3 -> Geometry m a -> Geometry m a
moving a b = Reified . Moving a b . geometryRep

-- Matrices (Geometry.hs:394-441): plain data, restated from their mathematical definitions.
translate :: Vec3 -> M44 Double
translate (V3 x y z) = V4 (V4 1 0 0 x) (V4 0 1 0 y) (V4 0 0 1 z) (V4 0 0 0 1)

rotateX, rotateY, rotateZ :: Double -> M44 Double
rotateX a = V4 (V4 1 0 0 0) (V4 0 c (-s) 0) (V4 0 s c 0) (V4 0 0 0 1) where (c, s) = (cos a, sin a)
rotateY a = V4 (V4 c 0 s 0) (V4 0 1 0 0) (V4 (-s) 0 c 0) (V4 0 0 0 1) where (c, s) = (cos a, sin a)
rotateZ a = V4 (V4 c (-s) 0 0) (V4 s c 0 0) (V4 0 0 1 0) (V4 0 0 0 1) where (c, s) = (cos a, sin a)

-- | Only legal through 'transformVertices' ('transform' needs a Euclidean matrix, Geometry.hs:379-381,434).
scale :: Double -> M44 Double
scale k = V4 (V4 k 0 0 0) (V4 0 k 0 0) (V4 0 0 k 0) (V4 0 0 0 1)

transformVertices :: M44 Double -> Mesh -> Mesh
transformVertices m (Mesh vs ts fs) = Mesh (A.compute (A.map f vs)) ts fs
  where f (V3 x y z) = let V4 a b c _ = m !* V4 x y z 1 in V3 a b c

-- | @.obj@ parsing is done by the library's loader (rt_obj_parse), which follows the rules and error strings of the
-- reference's parseObj (Geometry.hs:207-285; checked string for string in tests/test_host.py).
parseObj :: String -> Either String Mesh
parseObj text = unsafePerformIO $ withCStringLen text $ \(p, n) -> alloca $ \out -> do
  st <- c_obj_parse p (fromIntegral n) out
  if st /= rtOk then Left <$> (c_rt_last_error >>= peekCString) else do
    h <- peek out
    (nv, nt, nf) <- alloca $ \a -> alloca $ \b -> alloca $ \c -> do
      _ <- c_mesh_sizes h a b c
      (,,) <$> peek a <*> peek b <*> peek c
    (pv, pt, pf) <- alloca $ \a -> alloca $ \b -> alloca $ \c -> do
      _ <- c_mesh_data h a b c
      (,,) <$> peek a <*> peek b <*> peek c
    vs <- peekArray (3 * fromIntegral nv) pv
    ts <- peekArray (2 * fromIntegral nt) pt
    fs <- peekArray (6 * fromIntegral nf) pf
    _ <- c_mesh_destroy h
    let verts = A.fromList A.Seq [ V3 x y z | [x, y, z] <- chunks 3 vs ]
        uvs = A.fromList A.Seq [ V2 u v | [u, v] <- chunks 2 ts ]
        ix t = if t < 0 then Nothing else Just (fromIntegral t)
        faces = [ V3 (fromIntegral a, ix ta) (fromIntegral b, ix tb) (fromIntegral c, ix tc) | [a, ta, b, tb, c, tc] <- chunks 6 fs ]
    pure (Right (Mesh verts uvs faces))
  where chunks k xs = if null xs then [] else take k xs : chunks k (drop k xs)

readObj :: FilePath -> IO (Either String Mesh)
readObj path = parseObj <$> readFile path
