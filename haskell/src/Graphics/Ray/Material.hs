-- | Drop-in for @Graphics.Ray.Material@ (Material.hs:3-6).  The ten library materials are data; the raw 'Material'
-- constructor (a user closure) is kept so existing programs compile, and is rejected by 'raytrace' with an explicit
-- error.  'MaterialResult' is re-declared unchanged for source compatibility of user code that mentions it.
module Graphics.Ray.Material
  ( Material(..), MaterialResult(..)
  , lightSource, pitchBlack, lambertian, lommelSeeliger, mirror, metal, dielectric, transparent, isotropic, anisotropic
  ) where

import Graphics.Ray.Core (Color, HitRecord, Vec3)
import Graphics.Ray.Texture (Texture)
import Control.Monad.State (State)
import System.Random (StdGen)

data MaterialResult                       -- Material.hs:22-34
  = Absorb
  | Scatter Color Vec3
  | HemisphereF (Vec3 -> Color)
  | SphereF (Vec3 -> Color)

data Material
  = Material (Vec3 -> HitRecord -> (Color, State StdGen MaterialResult))   -- ^ user closure → RT_ERR_USER_DEFINED
  | LightSource Texture            -- ^ RT_MAT_LIGHT            Material.hs:41-42
  | PitchBlack                     -- ^ RT_MAT_PITCH_BLACK      Material.hs:46-47
  | Lambertian Texture             -- ^ RT_MAT_LAMBERTIAN       Material.hs:51-53
  | LommelSeeliger Texture         -- ^ RT_MAT_LOMMEL_SEELIGER  Material.hs:56-61
  | Mirror Texture                 -- ^ RT_MAT_MIRROR           Material.hs:64-67
  | Metal Double Texture           -- ^ RT_MAT_METAL            Material.hs:72-78
  | Dielectric Double              -- ^ RT_MAT_DIELECTRIC       Material.hs:89-106
  | Transparent Texture            -- ^ RT_MAT_TRANSPARENT      Material.hs:109-112
  | Isotropic Texture              -- ^ RT_MAT_ISOTROPIC        Material.hs:116-118
  | Anisotropic Double Texture     -- ^ RT_MAT_ANISOTROPIC      Material.hs:124-129

lightSource, lambertian, lommelSeeliger, mirror, transparent, isotropic :: Texture -> Material
lightSource = LightSource
lambertian = Lambertian
lommelSeeliger = LommelSeeliger
mirror = Mirror
transparent = Transparent
isotropic = Isotropic

pitchBlack :: Material
pitchBlack = PitchBlack

metal :: Double -> Texture -> Material
metal = Metal

dielectric :: Double -> Material
dielectric = Dielectric

anisotropic :: Double -> Texture -> Material
anisotropic = Anisotropic
